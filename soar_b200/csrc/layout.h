// Private layout of the three caller-owned scratch buffers (geom / binning / img).
// Plays the role of GeometryState / BinningState / ImageState
// (DGR/cuda_rasterizer/rasterizer_impl.h:22-84) but the layout is our own:
//   * one 96-byte, 32-byte-sector-aligned record per Gaussian holding exactly the
//     floats the blend kernels gather (the reference scatters them over 8 arrays
//     and carries 10 Jinv floats of which only 6 are ever read);
//   * keys/values double buffers for the onesweep passes + its histograms and
//     look-back status words;
//   * per-pixel final_T / n_contrib / final_D and per-tile ranges.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include "common.cuh"

namespace soar {

// ---- per-Gaussian record (24 floats = 96 B = 3 sectors) ------------------------
// float4 #0 and #1 are all a pixel needs to decide whether a Gaussian contributes
// (REC_THR: conservative lower bound on `power` below which alpha < 1/255 for sure).
constexpr int REC_FLOATS = 24;
enum RecSlot {
  REC_X = 0, REC_Y = 1, REC_CON_A = 2, REC_CON_B = 3,      // float4 #0
  REC_CON_C = 4, REC_OPAC = 5, REC_THR = 6, REC_DEPTH = 7, // float4 #1
  REC_R = 8, REC_G = 9, REC_B = 10, REC_J6 = 11,           // float4 #2
  REC_NX = 12, REC_NY = 13, REC_NZ = 14, REC_J9 = 15,      // float4 #3
  REC_J0 = 16, REC_J1 = 17, REC_J2 = 18, REC_J3 = 19,      // float4 #4
  REC_RECT_X = 20, REC_RECT_Y = 21, REC_RADIUS = 22, REC_VIEWCOS = 23  // float4 #5 (ints bit-cast)
};

// ---- decoupled look-back scan of tiles_touched ----------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// ---- onesweep radix sort ------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 12;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;  // 3072 pairs per CTA
// One tile size for every input: 1024-pair tiles for the small sorts (the depth sort of P = 500k
// Gaussians) were measured slower on the B200 (58 vs 50 us for the four passes, tools/sort_probe.cu):
// three times the tiles make the look-back chain and the status traffic longer than the shorter
// per-CTA work saves.
inline int sort_tiles_for(size_t n) { return (int)((n + SORT_TILE - 1) / SORT_TILE); }
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_MAX_PASSES = 4;    // 32-bit keys
constexpr int SORT_MAX_PASSES_R = 4;  // tile-id sort (up to 32 tile bits)
inline int sort_passes(int end_bit) { return (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS; }
// The key bits are spread evenly over the passes (12 tile bits sort as 6+6, not 8+4): the same number
// of ballots in total, but fewer digits to publish / look back over and longer scatter runs per digit.
__host__ __device__ inline int sort_bits_per_pass(int end_bit, int passes) {
  return passes > 0 ? (end_bit + passes - 1) / passes : 0;
}
__host__ __device__ inline int sort_pass_shift(int p, int end_bit, int passes) {
  return p * sort_bits_per_pass(end_bit, passes);
}
__host__ __device__ inline int sort_pass_bits(int p, int end_bit, int passes) {
  const int per = sort_bits_per_pass(end_bit, passes), left = end_bit - p * per;
  return left < per ? left : per;
}

struct GeomLayout {
  size_t records;        // float[P*24]
  size_t tiles_touched;  // u32[P]
  size_t point_offsets;  // u32[P] inclusive prefix sum of tiles_touched IN DEPTH ORDER
  size_t depth_keys[2];  // u32[P] x2: depth bits (0xffffffff for culled), ping-pong
  size_t perm[2];        // u32[P] x2: Gaussian indices, ping-pong; perm[0] holds the depth order
  size_t zero_begin, zero_bytes;  // region zeroed per call:
  size_t scan_status;    //   u64[n_scan_tiles] + u32 ticket + u32 error flag
  size_t scan_status_bytes;
  size_t ghist;          //   u32[4*256] digit histograms of the depth sort
  size_t sort_status;    //   u32[4*n_sort_tiles*256] look-back words + 4 tickets
  size_t total;
  int n_sort_tiles;
};

inline GeomLayout geom_layout(size_t P) {
  GeomLayout L;
  size_t off = 0;
  L.records = off;        off = align_up(off + P * REC_FLOATS * sizeof(float), 256);
  L.tiles_touched = off;  off = align_up(off + P * sizeof(uint32_t), 256);
  L.point_offsets = off;  off = align_up(off + P * sizeof(uint32_t), 256);
  for (int i = 0; i < 2; ++i) { L.depth_keys[i] = off; off = align_up(off + P * sizeof(uint32_t), 256); }
  for (int i = 0; i < 2; ++i) { L.perm[i] = off; off = align_up(off + P * sizeof(uint32_t), 256); }
  size_t n_tiles = (P + SCAN_TILE - 1) / SCAN_TILE;
  L.zero_begin = off;
  L.scan_status = off;
  L.scan_status_bytes = align_up(n_tiles * sizeof(uint64_t) + 16, 256);
  off += L.scan_status_bytes;
  L.ghist = off;          off = align_up(off + SORT_MAX_PASSES * SORT_RADIX * 4, 256);
  L.n_sort_tiles = sort_tiles_for(P);
  L.sort_status = off;
  off = align_up(off + ((size_t)SORT_MAX_PASSES * L.n_sort_tiles * SORT_RADIX + SORT_MAX_PASSES) * 4, 256);
  L.zero_bytes = off - L.zero_begin;
  L.total = off + 256;
  return L;
}

// ---- per-pixel / per-tile state ---------------------------------------------------
struct ImageLayout {
  size_t final_T;    // f32[W*H]
  size_t n_contrib;  // u32[W*H]
  size_t final_D;    // f32[W*H]
  size_t ranges;     // uint2[tiles] + two uint2: {non-empty tiles, forward tile counter}, {backward tile counter, -}
  size_t total;
};

inline ImageLayout image_layout(int W, int H) {
  ImageLayout L;
  size_t N = (size_t)W * H;
  size_t tiles = (size_t)ceil_div(W, TILE_X) * ceil_div(H, TILE_Y);
  size_t off = 0;
  L.final_T = off;   off = align_up(off + N * 4, 256);
  L.n_contrib = off; off = align_up(off + N * 4, 256);
  L.final_D = off;   off = align_up(off + N * 4, 256);
  L.ranges = off;    off = align_up(off + (tiles + 2) * 8, 256);
  L.total = off + 256;
  return L;
}

// ---- (tile id, Gaussian index) pairs ------------------------------------------------
struct BinningLayout {
  size_t keys[2];    // u32[R] x2 tile ids (ping-pong)
  size_t values[2];  // u32[R] x2 Gaussian indices
  size_t ghist;      // u32[passes*256] digit histograms -> exclusive offsets
  size_t status;     // u32[passes*n_sort_tiles*256] look-back words + u32 ticket per pass
  size_t zero_begin, zero_bytes;  // region that must be zero before sorting (ghist+status)
  size_t total;
  int n_sort_tiles;
};

inline BinningLayout binning_layout(size_t R, int passes) {
  BinningLayout L;
  size_t off = 0;
  L.keys[0] = off;   off = align_up(off + R * 4, 256);
  L.keys[1] = off;   off = align_up(off + R * 4, 256);
  L.values[0] = off; off = align_up(off + R * 4, 256);
  L.values[1] = off; off = align_up(off + R * 4, 256);
  L.n_sort_tiles = sort_tiles_for(R);
  L.zero_begin = off;
  L.ghist = off;     off = align_up(off + SORT_MAX_PASSES_R * SORT_RADIX * 4, 256);
  L.status = off;
  off = align_up(off + ((size_t)passes * L.n_sort_tiles * SORT_RADIX + SORT_MAX_PASSES_R) * 4, 256);
  L.zero_bytes = off - L.zero_begin;
  L.total = off + 256;
  return L;
}

// ---- backward scratch: per-Gaussian accumulator written by the blend backward ----
// 16 floats. The blend backward adds, per (warp, Gaussian) visit, the warp's sums of the
// colour / depth / opacity partials and of the six dL/ddist moments (sum of dd, dd*dx, dd*dy,
// dd*dx^2, dd*dx*dy, dd*dy^2 with d = centre - pixel); the per-Gaussian geometry backward turns
// the moments into dL/dmean2D and dL/dconic (backward.cu:820-855) once per Gaussian.
constexpr int GACC_FLOATS = 16;
enum GaccSlot {
  GA_DC0 = 0, GA_DC1 = 1, GA_DC2 = 2, GA_DD = 3, GA_DO = 4, GA_MX = 5, GA_MY = 6, GA_MXX = 7,
  GA_MXY = 8, GA_MYY = 9, GA_DN0 = 10, GA_DN1 = 11, GA_DN2 = 12, GA_GD = 13, GA_COUNT = 14
};

}  // namespace soar
