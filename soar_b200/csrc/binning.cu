// K2..K5 — binning for sm_100a:
//   * single-pass decoupled look-back inclusive scan of tiles_touched
//     (replaces cub::DeviceScan::InclusiveSum, DGR rasterizer_impl.cu:242-245);
//   * key duplication, key = tile_id << 32 | depth bits (rasterizer_impl.cu:66-99);
//   * in-house onesweep LSD radix sort of (u64 key, u32 value) pairs — one
//     histogram pass for all digit places, then one chained-scan scatter pass per
//     8-bit digit with per-digit decoupled look-back (replaces
//     cub::DeviceRadixSort::SortPairs[Descending], rasterizer_impl.cu:266-285);
//     stable, ascending or descending over bits [0, end_bit);
//   * tile ranges from the sorted keys (rasterizer_impl.cu:104-124).
// All of it is integer/byte work bound by HBM traffic; nothing here is shaped
// for tensor cores.
#include "common.cuh"
#include "layout.h"
#include "kernels.h"

namespace soar {

// =================================================================================
// Inclusive scan, decoupled look-back. status word: [63:62] flag, [31:0] value.
// =================================================================================
constexpr unsigned long long SCAN_FLAG_AGG = 1ull << 62;
constexpr unsigned long long SCAN_FLAG_INCL = 2ull << 62;
constexpr unsigned long long SCAN_FLAG_MASK = 3ull << 62;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n,
                     unsigned long long* status, uint32_t* ticket) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int base = tile * SCAN_TILE + tid * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  if (base + SCAN_ITEMS <= n) {
    const uint4* p = reinterpret_cast<const uint4*>(in + base);
    uint4 a = p[0], b = p[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < n) ? in[base + i] : 0u;
  }
  uint32_t tsum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) { tsum += v[i]; v[i] = tsum; }

  // block exclusive scan of thread sums
  uint32_t incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    uint32_t t = s_warp[w];
    if (w < warp) warp_off += t;
    block_total += t;
  }
  const uint32_t thread_excl = warp_off + incl - tsum;

  if (tid == 0) {
    uint32_t prefix = 0;
    if (tile == 0) {
      st_relaxed_u64(&status[0], SCAN_FLAG_INCL | block_total);
    } else {
      st_relaxed_u64(&status[tile], SCAN_FLAG_AGG | block_total);
      int t = (int)tile - 1;
      while (true) {
        unsigned long long s;
        do { s = ld_relaxed_u64(&status[t]); } while ((s & SCAN_FLAG_MASK) == 0);
        prefix += (uint32_t)s;
        if ((s & SCAN_FLAG_MASK) == SCAN_FLAG_INCL) break;
        --t;
      }
      st_relaxed_u64(&status[tile], SCAN_FLAG_INCL | (unsigned long long)(prefix + block_total));
    }
    s_prefix = prefix;
  }
  __syncthreads();
  const uint32_t add = s_prefix + thread_excl;
  if (base + SCAN_ITEMS <= n) {
    uint4 a = make_uint4(v[0] + add, v[1] + add, v[2] + add, v[3] + add);
    uint4 b = make_uint4(v[4] + add, v[5] + add, v[6] + add, v[7] + add);
    uint4* p = reinterpret_cast<uint4*>(out + base);
    p[0] = a; p[1] = b;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
      if (base + i < n) out[base + i] = v[i] + add;
  }
}

int launch_scan(const uint32_t* in, uint32_t* out, int n, void* status_region,
                size_t status_bytes, cudaStream_t stream) {
  if (n <= 0) return SOAR_OK;
  const int n_tiles = ceil_div(n, SCAN_TILE);
  // region layout: u64 status[n_tiles] | u32 ticket | u32 error flag (zeroed by caller)
  unsigned long long* status = reinterpret_cast<unsigned long long*>(status_region);
  uint32_t* ticket = reinterpret_cast<uint32_t*>(status + n_tiles);
  (void)status_bytes;
  {
    KernelScope ks(K_SCAN, stream);
    scan_lookback_kernel<<<n_tiles, SCAN_THREADS, 0, stream>>>(in, out, n, status, ticket);
  }
  SOAR_LAUNCH_CHECK("scan_lookback_kernel");
  return SOAR_OK;
}

// =================================================================================
// Key duplication
// =================================================================================
__global__ void __launch_bounds__(256)
duplicate_keys_kernel(int P, const float* __restrict__ records,
                      const uint32_t* __restrict__ tiles_touched,
                      const uint32_t* __restrict__ offsets, unsigned long long* __restrict__ keys,
                      uint32_t* __restrict__ values, uint32_t grid_x) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  if (tiles_touched[idx] == 0) return;
  uint32_t off = (idx == 0) ? 0u : offsets[idx - 1];
  const float4 r1 = reinterpret_cast<const float4*>(records + (size_t)idx * REC_FLOATS)[1];
  const float4 r5 = reinterpret_cast<const float4*>(records + (size_t)idx * REC_FLOATS)[5];
  const uint32_t depth_bits = __float_as_uint(r1.z);
  const uint32_t rx = __float_as_uint(r5.x), ry = __float_as_uint(r5.y);
  const uint32_t x0 = rx & 0xffffu, x1 = rx >> 16, y0 = ry & 0xffffu, y1 = ry >> 16;
  for (uint32_t y = y0; y < y1; ++y) {
    for (uint32_t x = x0; x < x1; ++x) {
      unsigned long long key = (unsigned long long)(y * grid_x + x);
      key <<= 32;
      key |= depth_bits;
      keys[off] = key;
      values[off] = (uint32_t)idx;
      ++off;
    }
  }
}

int launch_duplicate_keys(int P, const float* records, const uint32_t* tiles_touched,
                          const uint32_t* offsets, unsigned long long* keys, uint32_t* values,
                          uint32_t grid_x, cudaStream_t stream) {
  if (P <= 0) return SOAR_OK;
  {
    KernelScope ks(K_DUPLICATE, stream);
    duplicate_keys_kernel<<<ceil_div(P, 256), 256, 0, stream>>>(P, records, tiles_touched, offsets,
                                                                keys, values, grid_x);
  }
  SOAR_LAUNCH_CHECK("duplicate_keys_kernel");
  return SOAR_OK;
}

// =================================================================================
// Onesweep radix sort
// =================================================================================
constexpr uint32_t RS_FLAG_AGG = 1u << 30;
constexpr uint32_t RS_FLAG_INCL = 2u << 30;
constexpr uint32_t RS_FLAG_MASK = 3u << 30;
constexpr uint32_t RS_VALUE_MASK = (1u << 30) - 1u;

// One pass over the keys builds the digit histograms of every pass.
__global__ void __launch_bounds__(256)
radix_hist_kernel(const unsigned long long* __restrict__ keys, uint32_t n, int passes, int end_bit,
                  unsigned long long flip, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t sh[SORT_MAX_PASSES * SORT_RADIX];
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned long long k = keys[i] ^ flip;
    for (int p = 0; p < passes; ++p) {
      const int shift = p * SORT_RADIX_BITS;
      const int bits = min(SORT_RADIX_BITS, end_bit - shift);
      const uint32_t d = (uint32_t)(k >> shift) & ((1u << bits) - 1u);
      atomicAdd(&sh[p * SORT_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
    const uint32_t c = sh[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

// ghist[p][d] -> exclusive prefix over d, one CTA per pass.
__global__ void __launch_bounds__(SORT_RADIX) radix_hist_scan_kernel(uint32_t* ghist) {
  __shared__ uint32_t s_warp[SORT_RADIX / 32];
  uint32_t* h = ghist + blockIdx.x * SORT_RADIX;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t c = h[tid];
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t off = 0;
  for (int w = 0; w < warp; ++w) off += s_warp[w];
  h[tid] = off + incl - c;
}

__global__ void __launch_bounds__(SORT_THREADS)
radix_onesweep_kernel(const unsigned long long* __restrict__ keys_in,
                      unsigned long long* __restrict__ keys_out,
                      const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                      uint32_t n, int shift, uint32_t digit_mask, unsigned long long flip,
                      const uint32_t* __restrict__ ghist_excl, uint32_t* status,
                      uint32_t* ticket) {
  __shared__ unsigned long long s_keys[SORT_TILE];
  __shared__ uint32_t s_vals[SORT_TILE];
  __shared__ uint32_t s_warp_hist[SORT_THREADS / 32][SORT_RADIX];
  __shared__ uint32_t s_gofs[SORT_RADIX];
  __shared__ uint32_t s_digit_excl[SORT_RADIX];
  __shared__ uint32_t s_scan[SORT_THREADS / 32];
  __shared__ uint32_t s_tile;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NWARPS = SORT_THREADS / 32;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < NWARPS; ++w) s_warp_hist[w][tid] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * SORT_TILE;
  const uint32_t n_valid = min((uint32_t)SORT_TILE, n - tile_base);

  // ---- load (warp-striped inside each warp's contiguous chunk) and rank ----------
  unsigned long long key[SORT_ITEMS];
  uint32_t rank[SORT_ITEMS];
  const uint32_t warp_base = warp * (32 * SORT_ITEMS);
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; ++j) {
    const uint32_t local = warp_base + j * 32 + lane;
    key[j] = (local < n_valid) ? keys_in[tile_base + local] : (~0ull ^ flip);
  }
  const uint32_t lt_mask = lanemask_lt();
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; ++j) {
    const uint32_t d = (uint32_t)((key[j] ^ flip) >> shift) & digit_mask;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t lt = __popc(peers & lt_mask);
    const uint32_t pre = s_warp_hist[warp][d];
    __syncwarp();
    if (lt == 0) s_warp_hist[warp][d] = pre + __popc(peers);
    __syncwarp();
    rank[j] = pre + lt;
  }
  __syncthreads();

  // ---- per-digit: prefix over warps, publish aggregate, tile-local digit offsets --
  {
    const int d = tid;  // SORT_THREADS == SORT_RADIX
    uint32_t running = 0;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) {
      const uint32_t t = s_warp_hist[w][d];
      s_warp_hist[w][d] = running;
      running += t;
    }
    const uint32_t total = running;
    uint32_t* my_status = status + (size_t)tile * SORT_RADIX + d;
    if (tile == 0) st_relaxed_u32(my_status, RS_FLAG_INCL | total);
    else st_relaxed_u32(my_status, RS_FLAG_AGG | total);

    // block exclusive scan of `total` over digits
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_scan[w];
    const uint32_t digit_excl = off + incl - total;
    s_digit_excl[d] = digit_excl;

    // decoupled look-back over predecessor tiles for this digit
    uint32_t prev = 0;
    if (tile != 0) {
      int t = (int)tile - 1;
      while (true) {
        uint32_t s;
        const uint32_t* ps = status + (size_t)t * SORT_RADIX + d;
        do { s = ld_relaxed_u32(ps); } while ((s & RS_FLAG_MASK) == 0);
        prev += s & RS_VALUE_MASK;
        if ((s & RS_FLAG_MASK) == RS_FLAG_INCL) break;
        --t;
      }
      st_relaxed_u32(my_status, RS_FLAG_INCL | ((prev + total) & RS_VALUE_MASK));
    }
    s_gofs[d] = ghist_excl[d] + prev - digit_excl;
  }
  __syncthreads();

  // ---- exchange through shared memory so the global scatter is run-coalesced -----
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; ++j) {
    const uint32_t d = (uint32_t)((key[j] ^ flip) >> shift) & digit_mask;
    const uint32_t pos = s_digit_excl[d] + s_warp_hist[warp][d] + rank[j];
    rank[j] = pos;
    s_keys[pos] = key[j];
  }
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; ++j) {
    const uint32_t local = warp_base + j * 32 + lane;
    if (local < n_valid) s_vals[rank[j]] = vals_in[tile_base + local];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t p = i * SORT_THREADS + tid;
    if (p < n_valid) {
      const unsigned long long k = s_keys[p];
      const uint32_t d = (uint32_t)((k ^ flip) >> shift) & digit_mask;
      const uint32_t dst = s_gofs[d] + p;
      keys_out[dst] = k;
      vals_out[dst] = s_vals[p];
    }
  }
}

// Sort `n` pairs by key bits [0, end_bit). Input in keys[0]/values[0]; returns the
// index (0/1) of the buffer pair holding the result through *result_buf.
// `ghist` (SORT_MAX_PASSES*256 u32) and `status` must be zero on entry.
int launch_radix_sort(unsigned long long* keys[2], uint32_t* values[2], uint32_t n, int end_bit,
                      bool descending, uint32_t* ghist, uint32_t* status, int n_sort_tiles,
                      int* result_buf, cudaStream_t stream) {
  *result_buf = 0;
  if (n == 0 || end_bit <= 0) return SOAR_OK;
  const int passes = sort_passes(end_bit);
  if (passes > SORT_MAX_PASSES) {
    set_error("radix sort: %d bits need more than %d passes", end_bit, SORT_MAX_PASSES);
    return SOAR_E_INVALID;
  }
  if (n > RS_VALUE_MASK) {
    set_error("radix sort: %u pairs exceed the 2^30 look-back counter range", n);
    return SOAR_E_INVALID;
  }
  const unsigned long long flip = descending ? ~0ull : 0ull;
  const int tiles = ceil_div((int)n, SORT_TILE);
  if (tiles > n_sort_tiles) {
    set_error("radix sort: status region sized for %d tiles, need %d", n_sort_tiles, tiles);
    return SOAR_E_WORKSPACE;
  }
  const int hist_blocks = min(tiles * 2, 148 * 8);
  {
    KernelScope ks(K_RADIX_HIST, stream);
    radix_hist_kernel<<<hist_blocks, 256, 0, stream>>>(keys[0], n, passes, end_bit, flip, ghist);
  }
  SOAR_LAUNCH_CHECK("radix_hist_kernel");
  {
    KernelScope ks(K_RADIX_HIST_SCAN, stream);
    radix_hist_scan_kernel<<<passes, SORT_RADIX, 0, stream>>>(ghist);
  }
  SOAR_LAUNCH_CHECK("radix_hist_scan_kernel");
  uint32_t* tickets = status + (size_t)passes * n_sort_tiles * SORT_RADIX;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * SORT_RADIX_BITS;
    const int bits = min(SORT_RADIX_BITS, end_bit - shift);
    const uint32_t mask = (1u << bits) - 1u;
    {
      KernelScope ks(K_RADIX_ONESWEEP, stream);
      radix_onesweep_kernel<<<tiles, SORT_THREADS, 0, stream>>>(
          keys[p & 1], keys[(p + 1) & 1], values[p & 1], values[(p + 1) & 1], n, shift, mask, flip,
          ghist + p * SORT_RADIX, status + (size_t)p * n_sort_tiles * SORT_RADIX, tickets + p);
    }
    SOAR_LAUNCH_CHECK("radix_onesweep_kernel");
  }
  *result_buf = passes & 1;
  return SOAR_OK;
}

// =================================================================================
// Tile ranges
// =================================================================================
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t L, const unsigned long long* __restrict__ keys, uint2* ranges) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t cur = (uint32_t)(keys[idx] >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
  }
  if (idx == L - 1) ranges[cur].y = L;
}

int launch_tile_ranges(uint32_t L, const unsigned long long* keys, uint2* ranges, int n_tiles,
                       cudaStream_t stream) {
  SOAR_CUDA_TRY(cudaMemsetAsync(ranges, 0, (size_t)n_tiles * sizeof(uint2), stream));
  if (L == 0) return SOAR_OK;
  {
    KernelScope ks(K_TILE_RANGES, stream);
    tile_ranges_kernel<<<ceil_div((int)L, 256), 256, 0, stream>>>(L, keys, ranges);
  }
  SOAR_LAUNCH_CHECK("tile_ranges_kernel");
  return SOAR_OK;
}

}  // namespace soar
