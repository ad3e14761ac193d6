// Shared pieces of the tile blend kernels (forward and backward):
//   * the pixel <-> thread mapping: a 16x16 tile is eight 8x4 pixel blocks, one per warp
//     (an 8x4 block is touched by fewer Gaussians than a 16x2 row pair, so more warp
//     iterations can be skipped);
//   * a conservative "can this Gaussian reach alpha >= 1/255 anywhere in this pixel
//     rectangle" test, used to drop (Gaussian, tile) pairs at staging time and to give every
//     warp the list of Gaussians that can touch ITS 8x4 block. A pair that is dropped would
//     have failed the reference's `alpha < 1/255` test (forward.cu:545 / backward.cu:663) at
//     every pixel of the rectangle, so skipping it changes neither images, n_contrib nor
//     gradients; pairs that survive still take the reference's exact per-pixel tests.
#pragma once
#include "common.cuh"
#include "layout.h"

namespace soar {

constexpr int BLK_W = 8, BLK_H = 4;  // pixel block of one warp
constexpr int N_BLK = TILE_PIX / 32;  // 8 blocks per tile: 2 across, 4 down

__device__ __forceinline__ uint32_t pix_x_in_tile(uint32_t warp, uint32_t lane) {
  return (warp & 1u) * BLK_W + (lane & 7u);
}
__device__ __forceinline__ uint32_t pix_y_in_tile(uint32_t warp, uint32_t lane) {
  return (warp >> 1) * BLK_H + (lane >> 3);
}

// q(d) = A dx^2 + 2 B dx dy + C dy^2, d = Gaussian centre - pixel, power = -q/2. A box of pixels MAY hold a
// pixel that passes `power >= thr_exact` when the minimum of q over the box is small enough; thr (REC_THR)
// already sits 0.02 below the exact threshold and a further 1e-5 * (magnitude of the terms) absorbs the
// rounding of both the box evaluation and the kernels' own power polynomial. Non positive-definite
// conics and NaNs are never culled. Masks: bit = 2 * (y / 4) + x / 8, one per 8x4 pixel block of the tile.

// ---------------------------------------------------------------------------------------------
// The pair kernels' mask (blend_fwd2 / blend_bwd2: warp = 8x8 pixel block). Same conservative
// question per box, asked with less arithmetic:
//   * the minimum of the convex q over a box that does not contain its centre (d = 0) lies on a face
//     the centre can "see": the x-face nearest to 0 when 0 is outside the box's x-range, and likewise
//     in y. At most TWO edge minimisations per box instead of four;
//   * four 8x8 boxes instead of eight 8x4 ones (a warp blends an 8x8 block; its bit is written to
//     both of its 8x4 positions, so the scalar twins read the same bytes and only do a little
//     needless work);
//   * the rounding slack 1e-5 * mag is formed once from the tile box, which bounds every sub-box.
// The test is conservative (it keeps every entry that can pass the exact per-pixel test), so images,
// n_contrib and gradients do not depend on how tight it is; the packed kernels and their one-pixel twins
// use different granularities of it and must agree bit for bit in the forward
// (tests/test_blend_twins_gpu.py).
// ---------------------------------------------------------------------------------------------
// bounding box of the tile's pixels that can still change (see live_box_publish): the masks' first test
// runs against it instead of the whole tile, at no extra cost
struct LiveBox {
  float x0, y0, w, h;   // origin (pixel coordinates inside the tile) and extent (x1 - x0, y1 - y0)
};
struct BoxAxis {   // one axis of a box in d-space: [lo, hi], the face nearest to 0 and whether 0 is outside
  float lo, hi, e;
  bool vis;
};
__device__ __forceinline__ BoxAxis box_axis(float hi, float len) {
  BoxAxis a;
  a.hi = hi;
  a.lo = hi - len;
  a.vis = a.lo > 0.f || a.hi < 0.f;
  a.e = a.lo > 0.f ? a.lo : a.hi;
  return a;
}
// min of q over the box (x, y), +inf-free: returns 0 when the centre is inside
__device__ __forceinline__ float box_qmin(float A, float B2, float C, float nBoverC, float nBoverA,
                                          const BoxAxis& x, const BoxAxis& y) {
  // x-face at dx = x.e: q = A e^2 + (2B e + C t) t, t = clamp(-B/C e, y.lo, y.hi)
  const float ty = fminf(fmaxf(nBoverC * x.e, y.lo), y.hi);
  const float vx = A * x.e * x.e + (B2 * x.e + C * ty) * ty;
  const float tx = fminf(fmaxf(nBoverA * y.e, x.lo), x.hi);
  const float vy = C * y.e * y.e + (B2 * y.e + A * tx) * tx;
  const float big = 3.0e38f;
  const float q = fminf(x.vis ? vx : big, y.vis ? vy : big);
  return (x.vis || y.vis) ? q : 0.f;
}

// The live box of a tile. The forward kernels drop list entries at staging time that cannot reach
// alpha >= 1/255 at any pixel that is NOT yet done: a done pixel ignores every later entry, so such an entry
// changes nothing (and the backward, which only replays entries before each pixel's last contributor, does
// not need it either). On opaque surfels most of a silhouette tile is done after a few entries and the
// box shrinks to the strip of background pixels, which almost nothing behind the surface touches: the
// long tail of the list then costs one cheap rejection per entry instead of five box tests and a walk.
// Each warp reduces its not-done pixels to (x0, y0, x1, y1) with four redux instructions and leaves
// them in shared memory before the round's first barrier; every thread merges the warps' boxes after it.
__device__ __forceinline__ void live_box_publish(uint4* slot, uint32_t lane, uint32_t x, uint32_t ylo,
                                                 uint32_t yhi, bool any_live) {
  // lanes with no live pixel contribute the neutral elements
  const uint32_t x0 = __reduce_min_sync(0xffffffffu, any_live ? x : 255u);
  const uint32_t x1 = __reduce_max_sync(0xffffffffu, any_live ? x + 1u : 0u);   // exclusive: 0 = none
  const uint32_t y0 = __reduce_min_sync(0xffffffffu, any_live ? ylo : 255u);
  const uint32_t y1 = __reduce_max_sync(0xffffffffu, any_live ? yhi + 1u : 0u);
  if (lane == 0) *slot = make_uint4(x0, y0, x1, y1);
}
template <int NWARPS>
__device__ __forceinline__ LiveBox live_box_merge(const uint4* slots) {
  uint32_t x0 = 255u, y0 = 255u, x1 = 0u, y1 = 0u;
#pragma unroll
  for (int w = 0; w < NWARPS; ++w) {
    const uint4 b = slots[w];
    x0 = min(x0, b.x); y0 = min(y0, b.y); x1 = max(x1, b.z); y1 = max(y1, b.w);
  }
  LiveBox L;   // an empty box (the CTA is about to leave its loop) degenerates to the whole tile
  const bool none = x1 == 0u;
  L.x0 = none ? 0.f : (float)x0;
  L.y0 = none ? 0.f : (float)y0;
  L.w = none ? 15.f : (float)(x1 - 1u - x0);
  L.h = none ? 15.f : (float)(y1 - 1u - y0);
  return L;
}

// 8-bit mask, each 8x8 block's decision written to both of its 8x4 bits.
__device__ __forceinline__ uint32_t tile_block_mask_pairs(const float4 q0, const float4 q1, float tile_px,
                                                          float tile_py, const LiveBox live) {
  const float A = q0.z, B = q0.w, C = q1.x, thr = q1.z;
  const float det = A * C - B * B;
  if (!(A > 0.f && C > 0.f && det > 0.f)) return 0xffu;
  const float gx = q0.x - tile_px, gy = q0.y - tile_py;  // centre relative to the tile origin
  const float nBoverC = __fdividef(-B, C), nBoverA = __fdividef(-B, A), B2 = B + B;
  const float mx = fmaxf(fabsf(gx - 15.f), fabsf(gx)), my = fmaxf(fabsf(gy - 15.f), fabsf(gy));
  const float mag = A * mx * mx + C * my * my + fabsf(B2) * mx * my;
  const float lim = 2.f * (1e-5f * mag - thr);     // keep a box when qmin <= lim
  const BoxAxis X = box_axis(gx - live.x0, live.w), Y = box_axis(gy - live.y0, live.h);
  if (box_qmin(A, B2, C, nBoverC, nBoverA, X, Y) > lim) return 0u;
  const BoxAxis x0 = box_axis(gx, 7.f), x1 = box_axis(gx - 8.f, 7.f);
  const BoxAxis y0 = box_axis(gy, 7.f), y1 = box_axis(gy - 8.f, 7.f);
  uint32_t mask = 0;
  if (!(box_qmin(A, B2, C, nBoverC, nBoverA, x0, y0) > lim)) mask |= 0x05u;        // bits 0, 2
  if (!(box_qmin(A, B2, C, nBoverC, nBoverA, x1, y0) > lim)) mask |= 0x0au;        // bits 1, 3
  if (!(box_qmin(A, B2, C, nBoverC, nBoverA, x0, y1) > lim)) mask |= 0x50u;        // bits 4, 6
  if (!(box_qmin(A, B2, C, nBoverC, nBoverA, x1, y1) > lim)) mask |= 0xa0u;        // bits 5, 7
  return mask;
}

// The same test refined to the eight 8x4 blocks for the one-pixel-per-thread kernels: an 8x8 box that
// fails cannot hold an 8x4 box that passes, so only the halves of surviving 8x8 boxes are tested.
__device__ __forceinline__ uint32_t tile_block_mask_fast(const float4 q0, const float4 q1, float tile_px,
                                                         float tile_py, const LiveBox live) {
  const float A = q0.z, B = q0.w, C = q1.x, thr = q1.z;
  const float det = A * C - B * B;
  if (!(A > 0.f && C > 0.f && det > 0.f)) return 0xffu;
  const float gx = q0.x - tile_px, gy = q0.y - tile_py;
  const float nBoverC = __fdividef(-B, C), nBoverA = __fdividef(-B, A), B2 = B + B;
  const float mx = fmaxf(fabsf(gx - 15.f), fabsf(gx)), my = fmaxf(fabsf(gy - 15.f), fabsf(gy));
  const float mag = A * mx * mx + C * my * my + fabsf(B2) * mx * my;
  const float lim = 2.f * (1e-5f * mag - thr);
  const BoxAxis X = box_axis(gx - live.x0, live.w), Y = box_axis(gy - live.y0, live.h);
  if (box_qmin(A, B2, C, nBoverC, nBoverA, X, Y) > lim) return 0u;
  const BoxAxis x0 = box_axis(gx, 7.f), x1 = box_axis(gx - 8.f, 7.f);
  uint32_t mask = 0;
#pragma unroll
  for (int h = 0; h < 2; ++h) {          // the two 8-row halves of the tile
    const BoxAxis y8 = box_axis(gy - 8.f * h, 7.f);
    const bool k0 = !(box_qmin(A, B2, C, nBoverC, nBoverA, x0, y8) > lim);
    const bool k1 = !(box_qmin(A, B2, C, nBoverC, nBoverA, x1, y8) > lim);
    if (k0 || k1) {
      const BoxAxis ya = box_axis(gy - 8.f * h, 3.f), yb = box_axis(gy - 8.f * h - 4.f, 3.f);
      if (k0 && !(box_qmin(A, B2, C, nBoverC, nBoverA, x0, ya) > lim)) mask |= 1u << (4 * h);
      if (k1 && !(box_qmin(A, B2, C, nBoverC, nBoverA, x1, ya) > lim)) mask |= 2u << (4 * h);
      if (k0 && !(box_qmin(A, B2, C, nBoverC, nBoverA, x0, yb) > lim)) mask |= 4u << (4 * h);
      if (k1 && !(box_qmin(A, B2, C, nBoverC, nBoverA, x1, yb) > lim)) mask |= 8u << (4 * h);
    }
  }
  return mask;
}

}  // namespace soar
