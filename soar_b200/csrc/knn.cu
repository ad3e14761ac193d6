// distCUDA2 — mean squared distance to the 3 nearest other points — for sm_100a.
//
// Behavioural contract: KNN/simple_knn.cu:186-222 (SimpleKNN::knn). The search is
// exact there (seed bound from Morton neighbours, then every 1024-point box that
// can still hold a closer point is scanned), so any exact search returns the
// same three squared distances; they are evaluated with the reference's
// expression (d = other - self; d.x*d.x + d.y*d.y + d.z*d.z, KNN/simple_knn.cu:133-146)
// and averaged as (b0 + b1 + b2) / 3.0f (:183) so the floats agree bit for bit.
// Self is excluded by index, duplicates count as distance 0 (Q15).
//
// What is ours: no host round trips (the reference copies min and max back,
// :198,201, and allocates 7 temporaries per call); the caller provides one
// workspace. Points are gathered once into Morton order so a warp owns 32
// spatially adjacent points, and a three-level box hierarchy (32 / 1024 /
// 32768 points) replaces the flat scan of all 1024-point boxes: a box is opened
// only if some lane of the warp still needs it, and leaves cost 32 distance
// evaluations instead of 1024.
#include <float.h>
#include "common.cuh"
#include "layout.h"
#include "kernels.h"

namespace soar {

constexpr int KNN_LEAF = 32;
constexpr int KNN_FAN = 32;

struct Aabb {
  float lo[3];
  float hi[3];
};

struct KnnLayout {
  size_t bbox;      // u32[6] ordered-int encoded min/max
  size_t sorted;    // float4[P] (x, y, z, original index bits)
  size_t boxes[3];  // Aabb per level
  size_t sort;      // BinningLayout region
  size_t total;
  int n_boxes[3];
  BinningLayout BL;
};

inline KnnLayout knn_layout(size_t P) {
  KnnLayout L;
  size_t off = 0;
  L.bbox = off;   off = align_up(off + 6 * 4, 256);
  L.sorted = off; off = align_up(off + P * 16, 256);
  size_t n = P;
  for (int l = 0; l < 3; ++l) {
    n = (n + KNN_FAN - 1) / KNN_FAN;
    L.n_boxes[l] = (int)n;
    L.boxes[l] = off;
    off = align_up(off + n * sizeof(Aabb), 256);
  }
  L.BL = binning_layout(P, 4);
  L.sort = off;
  off += L.BL.total;
  L.total = off + 256;
  return L;
}

// order-preserving float <-> uint encoding for atomicMin/atomicMax
__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void knn_bbox_init_kernel(uint32_t* bbox) {
  if (threadIdx.x < 3) bbox[threadIdx.x] = 0xffffffffu;
  else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256)
knn_bbox_kernel(int P, const float* __restrict__ pts, uint32_t* bbox) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = pts[3 * (size_t)i + c];
      lo[c] = fminf(lo[c], v);
      hi[c] = fmaxf(hi[c], v);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      atomicMin(&bbox[c], f2ord(lo[c]));
      atomicMax(&bbox[3 + c], f2ord(hi[c]));
    }
  }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FFu;
  x = (x | (x << 8)) & 0x0300F00Fu;
  x = (x | (x << 4)) & 0x030C30C3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox,
                  uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t code = 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float lo = ord2f(bbox[c]), hi = ord2f(bbox[3 + c]);
    const float ext = hi - lo;
    float t = ext > 0.f ? (pts[3 * (size_t)i + c] - lo) / ext : 0.f;
    t = fminf(fmaxf(t, 0.f), 1.f);
    code |= spread10((uint32_t)(t * 1023.f)) << c;
  }
  keys[i] = code;
  vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256)
knn_gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                  float4* __restrict__ sorted) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const uint32_t src = order[i];
  sorted[i] = make_float4(pts[3 * (size_t)src], pts[3 * (size_t)src + 1], pts[3 * (size_t)src + 2],
                          __uint_as_float(src));
}

// One warp builds one box from up to 32 children (points or boxes).
template <bool FROM_POINTS>
__global__ void __launch_bounds__(256)
knn_boxes_kernel(int n_children, const float4* __restrict__ pts, const Aabb* __restrict__ child,
                 Aabb* __restrict__ out, int n_out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= n_out) return;
  const int c = warp * KNN_FAN + lane;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (c < n_children) {
    if (FROM_POINTS) {
      const float4 p = pts[c];
      lo[0] = hi[0] = p.x; lo[1] = hi[1] = p.y; lo[2] = hi[2] = p.z;
    } else {
      const Aabb b = child[c];
#pragma unroll
      for (int k = 0; k < 3; ++k) { lo[k] = b.lo[k]; hi[k] = b.hi[k]; }
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
    }
  }
  if (lane == 0) {
    Aabb b;
#pragma unroll
    for (int k = 0; k < 3; ++k) { b.lo[k] = lo[k]; b.hi[k] = hi[k]; }
    out[warp] = b;
  }
}

__device__ __forceinline__ float box_dist2(const Aabb& b, const float3& p) {
  const float dx = fmaxf(0.f, fmaxf(b.lo[0] - p.x, p.x - b.hi[0]));
  const float dy = fmaxf(0.f, fmaxf(b.lo[1] - p.y, p.y - b.hi[1]));
  const float dz = fmaxf(0.f, fmaxf(b.lo[2] - p.z, p.z - b.hi[2]));
  return dx * dx + dy * dy + dz * dz;
}

__device__ __forceinline__ void keep3(const float3& self, const float4& other, float* best) {
  const float3 d = {other.x - self.x, other.y - self.y, other.z - self.z};
  // insertion into the sorted triple as a min/max network (6 FMNMX instead of 3 compares + 6 selects;
  // the same three smallest values in the same order). The clamp keeps the compare chain's treatment of
  // NaN and infinite distances (never inserted: the triple starts at FLT_MAX).
  const float dist = fminf(d.x * d.x + d.y * d.y + d.z * d.z, FLT_MAX);
  const float t0 = fmaxf(best[0], dist);
  best[0] = fminf(best[0], dist);
  const float t1 = fmaxf(best[1], t0);
  best[1] = fminf(best[1], t0);
  best[2] = fminf(best[2], t1);
}

__global__ void __launch_bounds__(256)
knn_search_kernel(int P, const float4* __restrict__ sorted, const Aabb* __restrict__ box0,
                  const Aabb* __restrict__ box1, const Aabb* __restrict__ box2, int n0, int n1,
                  int n2, float* __restrict__ out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = idx < P;
  const float4 me4 = valid ? sorted[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
  const float3 me = {me4.x, me4.y, me4.z};

  // seed bound: the warp's own leaf (the 31 other points of this 32-point box)
  float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  const int leaf_begin = (idx >> 5) << 5;
  if (valid) {
    const int leaf_end = min(P, leaf_begin + KNN_LEAF);
    for (int i = leaf_begin; i < leaf_end; ++i) {
      if (i == idx) continue;
      keep3(me, sorted[i], best);
    }
  }
  // The traversal keeps the seed's three distances (the own leaf is not visited again) and walks
  // every level starting with the warp's own box, so the bound tightens on the nearest leaves
  // before the distant boxes are tested against it. Any order finds the same three smallest
  // distances. idx >> 5 is warp-uniform (a warp is one leaf).
  //
  // The 32 children of an opened box are first tested all at once, one child per lane, against the
  // WARP: the bounding box of its 32 query points and the largest of their current bounds. In
  // float arithmetic that distance is never above a lane's own box distance (subtraction, product
  // and sum are monotone and the warp box contains every query), so a child it rejects is rejected
  // by every lane's exact test too; only the survivors pay the 32 per-lane tests.
  const uint32_t lane = threadIdx.x & 31;
  Aabb wb;
  wb.lo[0] = wb.hi[0] = me.x; wb.lo[1] = wb.hi[1] = me.y; wb.lo[2] = wb.hi[2] = me.z;
  if (!valid) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { wb.lo[k] = FLT_MAX; wb.hi[k] = -FLT_MAX; }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      wb.lo[k] = fminf(wb.lo[k], __shfl_xor_sync(0xffffffffu, wb.lo[k], o));
      wb.hi[k] = fmaxf(wb.hi[k], __shfl_xor_sync(0xffffffffu, wb.hi[k], o));
    }
  }
  auto warp_bound = [&]() {
    float r = valid ? best[2] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
    return r;
  };
  auto children_mask = [&](const Aabb* boxes, int base, int cnt, float bound) {
    bool cand = false;
    if ((int)lane < cnt) {
      const Aabb b = boxes[base + lane];
      const float dx = fmaxf(0.f, fmaxf(b.lo[0] - wb.hi[0], wb.lo[0] - b.hi[0]));
      const float dy = fmaxf(0.f, fmaxf(b.lo[1] - wb.hi[1], wb.lo[1] - b.hi[1]));
      const float dz = fmaxf(0.f, fmaxf(b.lo[2] - wb.hi[2], wb.lo[2] - b.hi[2]));
      cand = !(dx * dx + dy * dy + dz * dz > bound);
    }
    return __ballot_sync(0xffffffffu, cand);
  };
  // next child of `mask`: the own box first (when it is among them), then ascending
  auto take = [](uint32_t& mask, int own_pos) {
    int t = (own_pos >= 0 && own_pos < 32 && ((mask >> own_pos) & 1u)) ? own_pos : __ffs(mask) - 1;
    mask &= ~(1u << t);
    return t;
  };
  float wr = warp_bound();
  const int own0 = idx >> 5, own1 = own0 >> 5, own2 = own1 >> 5;
  for (int c2 = 0; c2 < n2; c2 += 32) {
    uint32_t m2 = children_mask(box2, c2, min(32, n2 - c2), wr);
    while (m2) {
      const int b2 = c2 + take(m2, own2 - c2);
      bool want2 = false;
      if (valid) want2 = !(box_dist2(box2[b2], me) > best[2]);
      if (!__any_sync(0xffffffffu, want2)) continue;
      const int base1 = b2 * KNN_FAN;
      uint32_t m1 = children_mask(box1, base1, min(n1, base1 + KNN_FAN) - base1, wr);
      while (m1) {
        const int b1 = base1 + take(m1, b2 == own2 ? own1 - base1 : -1);
        bool want1 = false;
        if (want2) want1 = !(box_dist2(box1[b1], me) > best[2]);
        if (!__any_sync(0xffffffffu, want1)) continue;
        const int base0 = b1 * KNN_FAN;
        uint32_t m0 = children_mask(box0, base0, min(n0, base0 + KNN_FAN) - base0, wr);
        if (b1 == own1) m0 &= ~(1u << (own0 - base0));  // scanned as the seed
        bool scanned = false;
        while (m0) {
          const int b0 = base0 + take(m0, -1);
          bool want0 = false;
          if (want1) want0 = !(box_dist2(box0[b0], me) > best[2]);
          if (!__any_sync(0xffffffffu, want0)) continue;
          if (want0) {
            const int i_end = min(P, (b0 + 1) * KNN_LEAF);
            for (int i = b0 * KNN_LEAF; i < i_end; ++i) keep3(me, sorted[i], best);
          }
          scanned = true;
        }
        if (scanned) wr = warp_bound();  // the bounds only shrink: a stale value stays conservative
      }
    }
  }
  if (valid) out[__float_as_uint(me4.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

}  // namespace soar

using namespace soar;

extern "C" {

size_t soar_knn_workspace_bytes(int32_t P) { return knn_layout(P < 0 ? 0 : (size_t)P).total; }

int soar_knn_dist2(const float* points, int32_t P, float* mean_dist2, void* workspace,
                   size_t workspace_bytes, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) { set_error("soar_knn_dist2: P < 0"); return SOAR_E_INVALID; }
  if (P == 0) return SOAR_OK;
  if (!points || !mean_dist2) { set_error("soar_knn_dist2: NULL pointer"); return SOAR_E_INVALID; }
  const KnnLayout L = knn_layout(P);
  if (!workspace || workspace_bytes < L.total) {
    set_error("soar_knn_dist2: workspace holds %zu bytes, need %zu", workspace_bytes, L.total);
    return SOAR_E_WORKSPACE;
  }
  char* ws = reinterpret_cast<char*>(workspace);
  uint32_t* bbox = reinterpret_cast<uint32_t*>(ws + L.bbox);
  float4* sorted = reinterpret_cast<float4*>(ws + L.sorted);
  Aabb* box[3] = {reinterpret_cast<Aabb*>(ws + L.boxes[0]), reinterpret_cast<Aabb*>(ws + L.boxes[1]),
                  reinterpret_cast<Aabb*>(ws + L.boxes[2])};
  char* sw = ws + L.sort;
  uint32_t* keys[2] = {reinterpret_cast<uint32_t*>(sw + L.BL.keys[0]),
                       reinterpret_cast<uint32_t*>(sw + L.BL.keys[1])};
  uint32_t* vals[2] = {reinterpret_cast<uint32_t*>(sw + L.BL.values[0]),
                       reinterpret_cast<uint32_t*>(sw + L.BL.values[1])};

  {
    KernelScope ks(K_KNN_BBOX, stream);
    knn_bbox_init_kernel<<<1, 32, 0, stream>>>(bbox);
  }
  const int red_blocks = min(ceil_div(P, 256), 148 * 8);
  {
    KernelScope ks(K_KNN_BBOX, stream);
    knn_bbox_kernel<<<red_blocks, 256, 0, stream>>>(P, points, bbox);
  }
  {
    KernelScope ks(K_KNN_MORTON, stream);
    knn_morton_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, points, bbox, keys[0], vals[0]);
  }
  SOAR_LAUNCH_CHECK("knn_morton_kernel");
  SOAR_CUDA_TRY(cudaMemsetAsync(sw + L.BL.zero_begin, 0, L.BL.zero_bytes, stream));
  int res = 0;
  int rc = launch_radix_sort(keys, vals, (uint32_t)P, 30, false, /*iota_values=*/false,
                             /*have_hist=*/false, reinterpret_cast<uint32_t*>(sw + L.BL.ghist),
                             reinterpret_cast<uint32_t*>(sw + L.BL.status), L.BL.n_sort_tiles, &res,
                             stream);
  if (rc) return rc;
  {
    KernelScope ks(K_KNN_GATHER, stream);
    knn_gather_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, points, vals[res], sorted);
  }
  {
    KernelScope ks(K_KNN_BOXES, stream);
    knn_boxes_kernel<true><<<ceil_div(L.n_boxes[0] * 32, 256), 256, 0, stream>>>(
        P, sorted, nullptr, box[0], L.n_boxes[0]);
  }
  {
    KernelScope ks(K_KNN_BOXES, stream);
    knn_boxes_kernel<false><<<ceil_div(L.n_boxes[1] * 32, 256), 256, 0, stream>>>(
        L.n_boxes[0], nullptr, box[0], box[1], L.n_boxes[1]);
  }
  {
    KernelScope ks(K_KNN_BOXES, stream);
    knn_boxes_kernel<false><<<ceil_div(L.n_boxes[2] * 32, 256), 256, 0, stream>>>(
        L.n_boxes[1], nullptr, box[1], box[2], L.n_boxes[2]);
  }
  {
    KernelScope ks(K_KNN_SEARCH, stream);
    knn_search_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, sorted, box[0], box[1], box[2],
                                                            L.n_boxes[0], L.n_boxes[1],
                                                            L.n_boxes[2], mean_dist2);
  }
  SOAR_LAUNCH_CHECK("knn_search_kernel");
  return SOAR_OK;
}

}  // extern "C"
