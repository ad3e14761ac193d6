// K2..K5 — binning for sm_100a.
//
// The reference (DGR/cuda_rasterizer/rasterizer_impl.cu:242-295) scans
// tiles_touched, writes one 64-bit key (tile << 32 | depth bits) per (Gaussian,
// tile) pair in Gaussian order, radix-sorts all R pairs over 32+b bits (6 byte
// passes at 1024^2) and finds the tile ranges. The sorted order it defines is
//      (tile, depth bits, Gaussian index)           [stable LSD sort, unique
//                                                    (Gaussian, tile) pairs]
// and that order — not the way it is reached — is the contract we are compared
// against bit for bit. Here it is reached with far less traffic:
//   1. the P Gaussians are sorted by depth bits ONCE (stable, so ties stay in
//      index order): 4 byte passes over P 8-byte pairs instead of over R pairs;
//   2. a decoupled look-back scan of tiles_touched in that order gives every
//      Gaussian its slot range;
//   3. pairs are emitted in depth order by warps that write whole contiguous
//      runs (coalesced), with a 32-bit key that is just the tile id; the digit
//      histograms of the tile sort are accumulated on the fly;
//   4. a stable onesweep sort over the b tile bits only (2 byte passes at
//      1024^2) of 8-byte pairs puts the pairs in (tile, depth, index) order;
//   5. tile ranges are read off the sorted tile ids.
// Descending order (SOAR's back-normal view, sort_descending) flips both sorts.
// All of it is integer/byte work bound by HBM traffic; nothing here is shaped
// for tensor cores.
#include "common.cuh"
#include "layout.h"
#include "kernels.h"

namespace soar {

// =================================================================================
// Inclusive scan, decoupled look-back. status word: [63:62] flag, [31:0] value.
// Optional gather: element k is in[index[k]].
// =================================================================================
constexpr unsigned long long SCAN_FLAG_AGG = 1ull << 62;
constexpr unsigned long long SCAN_FLAG_INCL = 2ull << 62;
constexpr unsigned long long SCAN_FLAG_MASK = 3ull << 62;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_lookback_kernel(const uint32_t* __restrict__ in, const uint32_t* __restrict__ index,
                     uint32_t* __restrict__ out, int n, unsigned long long* status,
                     uint32_t* ticket) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_prefix;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int base = tile * SCAN_TILE + tid * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  if (index != nullptr) {
    if (base + SCAN_ITEMS <= n) {
      const uint4* p = reinterpret_cast<const uint4*>(index + base);
      const uint4 a = p[0], b = p[1];
      v[0] = in[a.x]; v[1] = in[a.y]; v[2] = in[a.z]; v[3] = in[a.w];
      v[4] = in[b.x]; v[5] = in[b.y]; v[6] = in[b.z]; v[7] = in[b.w];
    } else {
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; ++i) v[i] = (base + i < n) ? in[index[base + i]] : 0u;
    }
  } else if (base + SCAN_ITEMS <= n) {
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

  // warp 0 resolves the tile's exclusive prefix: 32 predecessor status words per step
  if (warp == 0) {
    uint32_t prefix = 0;
    if (tile == 0) {
      if (lane == 0) st_relaxed_u64(&status[0], SCAN_FLAG_INCL | block_total);
    } else {
      if (lane == 0) st_relaxed_u64(&status[tile], SCAN_FLAG_AGG | block_total);
      int t = (int)tile - 1;
      while (true) {
        const int tt = t - lane;
        const unsigned long long s = tt >= 0 ? ld_relaxed_u64(&status[tt]) : SCAN_FLAG_INCL;
        const unsigned long long flag = s & SCAN_FLAG_MASK;
        const uint32_t not_ready = __ballot_sync(0xffffffffu, flag == 0);
        const uint32_t incl = __ballot_sync(0xffffffffu, flag == SCAN_FLAG_INCL);
        const uint32_t first_nr = not_ready ? (uint32_t)__ffs(not_ready) - 1u : 32u;
        const uint32_t first_incl = incl ? (uint32_t)__ffs(incl) - 1u : 32u;
        const bool finish = first_incl < first_nr;
        const uint32_t take = finish ? first_incl + 1u : first_nr;  // lanes [0, take)
        uint32_t val = (uint32_t)lane < take ? (uint32_t)s : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
        prefix += val;
        if (finish) break;
        t -= (int)take;
      }
      if (lane == 0)
        st_relaxed_u64(&status[tile], SCAN_FLAG_INCL | (unsigned long long)(prefix + block_total));
    }
    if (lane == 0) s_prefix = prefix;
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

int launch_scan(const uint32_t* in, const uint32_t* index, uint32_t* out, int n,
                void* status_region, cudaStream_t stream) {
  if (n <= 0) return SOAR_OK;
  const int n_tiles = ceil_div(n, SCAN_TILE);
  // region layout: u64 status[n_tiles] | u32 ticket | u32 error flag (zeroed by caller)
  unsigned long long* status = reinterpret_cast<unsigned long long*>(status_region);
  uint32_t* ticket = reinterpret_cast<uint32_t*>(status + n_tiles);
  {
    KernelScope ks(K_SCAN, stream);
    scan_lookback_kernel<<<n_tiles, SCAN_THREADS, 0, stream>>>(in, index, out, n, status, ticket);
  }
  SOAR_LAUNCH_CHECK("scan_lookback_kernel");
  return SOAR_OK;
}

// =================================================================================
// Pair emission in depth order. One warp owns 32 consecutive Gaussians of the
// depth-sorted permutation; their slot ranges are contiguous, so the warp writes
// one contiguous run of (tile id, Gaussian index) pairs, 32 at a time, each lane
// finding the owner of its slot with a 5-step search over the lanes' start
// offsets. Row-major over the tile rect like rasterizer_impl.cu:88-97.
// The digit histograms of the tile sort are accumulated here (no extra pass).
// =================================================================================
constexpr int DUP_THREADS = 256;

__global__ void __launch_bounds__(DUP_THREADS)
duplicate_tiles_kernel(int P, const float* __restrict__ records, const uint32_t* __restrict__ perm,
                       const uint32_t* __restrict__ offsets_incl, uint32_t* __restrict__ keys,
                       uint32_t* __restrict__ values, uint32_t grid_x, uint32_t flip, int passes,
                       int end_bit, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t s_hist[SORT_MAX_PASSES_R * SORT_RADIX];
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += DUP_THREADS) s_hist[i] = 0;
  __syncthreads();
  const int per_pass = sort_bits_per_pass(end_bit, passes);  // one division per thread, not per pair
  const uint32_t low_mask = (1u << min(per_pass, end_bit)) - 1u;
  uint32_t run_digit[SORT_MAX_PASSES_R - 1], run_count[SORT_MAX_PASSES_R - 1];
#pragma unroll
  for (int p = 0; p < SORT_MAX_PASSES_R - 1; ++p) { run_digit[p] = 0; run_count[p] = 0; }

  const uint32_t lane = lane_id();
  const int k = blockIdx.x * DUP_THREADS + threadIdx.x;
  uint32_t idx = 0, start = 0, tt = 0, xy0 = 0, w = 1;
  if (k < P) {
    idx = perm[k];
    const uint32_t incl = offsets_incl[k];
    const uint32_t prev = (k == 0) ? 0u : offsets_incl[k - 1];
    tt = incl - prev;
    start = prev;
    if (tt != 0) {
      const float4 r5 = __ldg(reinterpret_cast<const float4*>(records + (size_t)idx * REC_FLOATS) + 5);
      const uint32_t rx = __float_as_uint(r5.x), ry = __float_as_uint(r5.y);
      xy0 = (rx & 0xffffu) | (ry << 16);
      w = (rx >> 16) - (rx & 0xffffu);
    }
  } else {
    // pad lanes: start at the end of the warp's run so the search never selects them early
    start = 0xffffffffu;
  }
  const uint32_t magic = w > 1 ? 0xffffffffu / w + 1u : 0u;  // floor(2^32 / w) + 1 (w == 1: row = r)
  const uint32_t run_begin = __shfl_sync(0xffffffffu, start, 0);
  // end of the run = start + tt of the last valid lane
  uint32_t endv = (k < P) ? start + tt : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) endv = max(endv, __shfl_xor_sync(0xffffffffu, endv, o));
  const uint32_t run_end = endv;

  for (uint32_t slot0 = run_begin; slot0 < run_end; slot0 += 32) {
    const uint32_t slot = slot0 + lane;
    // owner = last lane whose start <= slot (zero-size lanes are skipped by "last")
    uint32_t j = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const uint32_t cand = j + step;
      const uint32_t v = __shfl_sync(0xffffffffu, start, cand & 31);
      if (v <= slot) j = cand;
    }
    const uint32_t o_start = __shfl_sync(0xffffffffu, start, j);
    const uint32_t o_xy0 = __shfl_sync(0xffffffffu, xy0, j);
    const uint32_t o_w = __shfl_sync(0xffffffffu, w, j);
    const uint32_t o_magic = __shfl_sync(0xffffffffu, magic, j);
    const uint32_t o_idx = __shfl_sync(0xffffffffu, idx, j);
    if (slot < run_end) {
      const uint32_t r = slot - o_start;
      // r / o_w with the owner's reciprocal: the estimate is exact or one too large
      uint32_t row = o_w > 1 ? __umulhi(r, o_magic) : r;
      if (row * o_w > r) --row;
      const uint32_t col = r - row * o_w;
      const uint32_t tile = ((o_xy0 >> 16) + row) * grid_x + (o_xy0 & 0xffffu) + col;
      keys[slot] = tile;
      values[slot] = o_idx;
      // digit histograms of the tile sort. A warp's 32 slots are consecutive tiles of a few rects:
      // their low digits are all different (conflict-free shared atomics); their higher digits are
      // nearly all equal, across the lanes (a 32-way same-address conflict) and from one slot of a
      // lane to its next, so every lane run-length encodes them and adds a whole run at once.
      const uint32_t kf = tile ^ flip;
      atomicAdd(&s_hist[kf & low_mask], 1u);
#pragma unroll
      for (int p = 1; p < SORT_MAX_PASSES_R; ++p) {
        if (p < passes) {
          const int shift = p * per_pass;
          const uint32_t d = (kf >> shift) & ((1u << min(per_pass, end_bit - shift)) - 1u);
          if (d != run_digit[p - 1]) {
            if (run_count[p - 1]) atomicAdd(&s_hist[p * SORT_RADIX + run_digit[p - 1]], run_count[p - 1]);
            run_digit[p - 1] = d;
            run_count[p - 1] = 0;
          }
          ++run_count[p - 1];
        }
      }
    }
  }
#pragma unroll
  for (int p = 1; p < SORT_MAX_PASSES_R; ++p)
    if (p < passes && run_count[p - 1])
      atomicAdd(&s_hist[p * SORT_RADIX + run_digit[p - 1]], run_count[p - 1]);
  __syncthreads();
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += DUP_THREADS) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

int launch_duplicate_tiles(int P, const float* records, const uint32_t* perm,
                           const uint32_t* offsets_incl, uint32_t* keys, uint32_t* values,
                           uint32_t grid_x, bool descending, int tile_bits, uint32_t* ghist,
                           cudaStream_t stream) {
  if (P <= 0) return SOAR_OK;
  const int passes = sort_passes(tile_bits);
  if (passes > SORT_MAX_PASSES_R) {
    set_error("duplicate: %d tile bits need more than %d passes", tile_bits, SORT_MAX_PASSES_R);
    return SOAR_E_INVALID;
  }
  {
    KernelScope ks(K_DUPLICATE, stream);
    duplicate_tiles_kernel<<<ceil_div(P, DUP_THREADS), DUP_THREADS, 0, stream>>>(
        P, records, perm, offsets_incl, keys, values, grid_x, descending ? 0xffffffffu : 0u, passes,
        tile_bits, ghist);
  }
  SOAR_LAUNCH_CHECK("duplicate_tiles_kernel");
  return SOAR_OK;
}

// =================================================================================
// Onesweep radix sort of (u32 key, u32 value) pairs
// =================================================================================
constexpr uint32_t RS_FLAG_AGG = 1u << 30;
constexpr uint32_t RS_FLAG_INCL = 2u << 30;
constexpr uint32_t RS_FLAG_MASK = 3u << 30;
constexpr uint32_t RS_VALUE_MASK = (1u << 30) - 1u;
// tools/sort_probe.cu compiles this file with SOAR_SORT_PROBE to ablate stages of the pass
#ifdef SOAR_SORT_PROBE
__constant__ int g_probe;
#define SORT_PROBE(bit) ((g_probe >> (bit)) & 1)
#else
#define SORT_PROBE(bit) 0
#endif
constexpr int RS_LB = 8;  // look-back words in flight per step
constexpr size_t sort_smem_bytes(int items) {
  return (2 * (size_t)SORT_THREADS * items + (SORT_THREADS / 32) * SORT_RADIX + 2 * SORT_RADIX +
          2 * (SORT_THREADS / 32) + 4) * 4;
}

// One pass over the keys builds the digit histograms of every pass.
__global__ void __launch_bounds__(256)
radix_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, int passes, int end_bit,
                  uint32_t flip, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t sh[SORT_MAX_PASSES * SORT_RADIX];
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  const int per_pass = sort_bits_per_pass(end_bit, passes);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t k = keys[i] ^ flip;
    for (int p = 0; p < passes; ++p) {
      const int shift = p * per_pass;
      const int bits = min(per_pass, end_bit - shift);
      const uint32_t d = (k >> shift) & ((1u << bits) - 1u);
      atomicAdd(&sh[p * SORT_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * SORT_RADIX; i += blockDim.x) {
    const uint32_t c = sh[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

// peers & (lanes whose digit has the same value in the bit `mask` as mine): bit test with a
// predicate result, ballot, select, one three-input logic op — four instructions per digit bit
// (the C++ spelling compiles to six).
__device__ __forceinline__ uint32_t peers_with_same_bit(uint32_t peers, uint32_t d, uint32_t mask) {
  uint32_t out;
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " .reg .b32 t, bal, m;\n"
      " and.b32 t, %2, %3;\n"
      " setp.ne.u32 p, t, 0;\n"
      " vote.sync.ballot.b32 bal, p, 0xffffffff;\n"
      " selp.b32 m, 0, 0xffffffff, p;\n"
      " lop3.b32 %0, %1, bal, m, 0x60;\n"
      "}"
      : "=r"(out)
      : "r"(peers), "r"(d), "r"(mask));
  return out;
}

// vals_in == nullptr: the value of element i is i (first pass of the depth sort).
// NBITS = digit width known at compile time (the ranking loop is then exactly NBITS ballots with no
// per-bit branch); launch_radix_sort picks the instantiation.
template <int NBITS, int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, 4)
radix_onesweep_kernel(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out,
                      const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                      uint32_t n, int shift, uint32_t flip,
                      const uint32_t* __restrict__ ghist_count, uint32_t* status, uint32_t* ticket) {
  constexpr uint32_t digit_mask = (1u << NBITS) - 1u;
  constexpr int TILE = SORT_THREADS * ITEMS;  // pairs per CTA
  // dynamic shared memory (> 48 KB): keys | vals | per-warp digit histograms | offsets
  extern __shared__ uint32_t s_dyn[];
  uint32_t* const s_keys = s_dyn;
  uint32_t* const s_vals = s_keys + TILE;
  uint32_t (*const s_warp_hist)[SORT_RADIX] =
      reinterpret_cast<uint32_t (*)[SORT_RADIX]>(s_vals + TILE);
  uint32_t* const s_gofs = s_vals + TILE + (SORT_THREADS / 32) * SORT_RADIX;
  uint32_t* const s_digit_excl = s_gofs + SORT_RADIX;
  uint32_t* const s_scan = s_digit_excl + SORT_RADIX;
  uint32_t* const s_gscan = s_scan + SORT_THREADS / 32;
  uint32_t& s_tile = s_gscan[SORT_THREADS / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NWARPS = SORT_THREADS / 32;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < NWARPS; ++w) s_warp_hist[w][tid] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * TILE;
  const uint32_t n_valid = min((uint32_t)TILE, n - tile_base);

  // ---- load (warp-striped inside each warp's contiguous chunk) and rank ----------
  uint32_t key[ITEMS];
  uint32_t val[ITEMS];
  uint32_t rank[ITEMS];
  const uint32_t warp_base = warp * (32 * ITEMS);
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t local = warp_base + j * 32 + lane;
    key[j] = (local < n_valid) ? keys_in[tile_base + local] : (~0u ^ flip);
  }
  // values are fetched now so their latency hides behind the ranking
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t local = warp_base + j * 32 + lane;
    val[j] = tile_base + local;
    if (vals_in != nullptr && local < n_valid) val[j] = vals_in[tile_base + local];
  }
  const uint32_t my_gcount = ghist_count[tid];  // digit tid (SORT_THREADS == SORT_RADIX)
  const uint32_t lt_mask = lanemask_lt();
  // Rank of a key among the warp's keys with the same digit = (count in earlier items) + (lanes
  // below in this item). The per-digit running count lives in shared memory; the first lane of
  // each peer group bumps it with ONE atomic whose return value is broadcast, so the twelve
  // items only depend on each other through same-address atomic ordering, not through
  // load/sync/store round trips.
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t d = ((key[j] ^ flip) >> shift) & digit_mask;
    if (SORT_PROBE(1)) { rank[j] = j; continue; }
    // lanes holding the same digit: one ballot per digit bit (constant time; match.any
    // serialises over the distinct values in the warp)
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < NBITS; ++b) {
      peers = peers_with_same_bit(peers, d, 1u << b);
    }
    const uint32_t lt = __popc(peers & lt_mask);
    const int leader = __ffs(peers) - 1;
    uint32_t pre = 0;
    if (lt == 0) pre = atomicAdd(&s_warp_hist[warp][d], (uint32_t)__popc(peers));
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[j] = pre + lt;
  }
  __syncthreads();

  // ---- per-digit: prefix over warps, publish aggregate, tile-local digit offsets --
  // (running the look-back after the shared-memory exchange instead was measured: faster on random
  // keys, 15 % slower on the real tile keys — the look-back stays here)
  const int dg = tid;  // this thread's digit (SORT_THREADS == SORT_RADIX)
  uint32_t dg_total, dg_excl, dg_ghist;
  uint32_t* my_status;
  bool live_digit;
  {
    const int d = dg;
    uint32_t running = 0;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) {
      const uint32_t t = s_warp_hist[w][d];
      s_warp_hist[w][d] = running;
      running += t;
    }
    const uint32_t total = running;
    // digits beyond the pass's width never occur: nothing to publish or to look back over
    live_digit = d < (1 << NBITS);
    my_status = status + (size_t)tile * SORT_RADIX + d;
    if (live_digit) {
      if (tile == 0) st_relaxed_u32(my_status, RS_FLAG_INCL | total);
      else st_relaxed_u32(my_status, RS_FLAG_AGG | total);
    }

    // block exclusive scans over digits of `total` (tile-local digit offsets) and of the global
    // digit counts (where each digit's output region starts; every CTA redoes this 256-entry scan
    // rather than paying a separate launch per sort for it)
    uint32_t incl = total, gincl = my_gcount;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      const uint32_t g = __shfl_up_sync(0xffffffffu, gincl, o);
      if (lane >= o) { incl += t; gincl += g; }
    }
    if (lane == 31) { s_scan[warp] = incl; s_gscan[warp] = gincl; }
    __syncthreads();
    uint32_t off = 0, goff = 0;
    for (int w = 0; w < warp; ++w) { off += s_scan[w]; goff += s_gscan[w]; }
    const uint32_t digit_excl = off + incl - total;
    const uint32_t my_ghist = goff + gincl - my_gcount;
    s_digit_excl[d] = digit_excl;
    dg_total = total; dg_excl = digit_excl; dg_ghist = my_ghist;
  }
  {
    const int d = dg;
    const uint32_t total = dg_total, digit_excl = dg_excl, my_ghist = dg_ghist;
    // decoupled look-back over predecessor tiles for this digit; RS_LB status words are
    // fetched per step so the chain costs one L2 round trip per RS_LB tiles, not per tile
    uint32_t prev = 0;
    if (tile != 0 && live_digit && !SORT_PROBE(0)) {
      int t = (int)tile - 1;
      bool found = false;
      while (!found) {
        uint32_t s[RS_LB];
#pragma unroll
        for (int k = 0; k < RS_LB; ++k) {
          const int tt = t - k;
          s[k] = tt >= 0 ? ld_relaxed_u32(status + (size_t)tt * SORT_RADIX + d) : RS_FLAG_INCL;
        }
        int used = 0;
#pragma unroll
        for (int k = 0; k < RS_LB; ++k) {
          if (!found && used == k && (s[k] & RS_FLAG_MASK) != 0) {
            prev += s[k] & RS_VALUE_MASK;
            used = k + 1;
            found = (s[k] & RS_FLAG_MASK) == RS_FLAG_INCL;
          }
        }
        t -= used;
      }
      st_relaxed_u32(my_status, RS_FLAG_INCL | ((prev + total) & RS_VALUE_MASK));
    }
    s_gofs[d] = my_ghist + prev - digit_excl;
  }
  __syncthreads();
  // ---- exchange through shared memory so the global scatter is run-coalesced -----
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t d = ((key[j] ^ flip) >> shift) & digit_mask;
    const uint32_t pos = s_digit_excl[d] + s_warp_hist[warp][d] + rank[j];
    rank[j] = pos;
    s_keys[pos] = key[j];
  }
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t local = warp_base + j * 32 + lane;
    if (local < n_valid) s_vals[rank[j]] = val[j];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const uint32_t p = i * SORT_THREADS + tid;
    if (p < n_valid) {
      const uint32_t k = s_keys[p];
      const uint32_t d = ((k ^ flip) >> shift) & digit_mask;
      const uint32_t dst = s_gofs[d] + p;
      keys_out[dst] = k;
      vals_out[dst] = s_vals[p];
    }
  }
}

typedef void (*OnesweepFn)(const uint32_t*, uint32_t*, const uint32_t*, uint32_t*, uint32_t, int,
                           uint32_t, const uint32_t*, uint32_t*, uint32_t*);
template <int ITEMS>
static OnesweepFn onesweep_for_bits(int bits) {
  switch (bits) {
    case 1: return radix_onesweep_kernel<1, ITEMS>;
    case 2: return radix_onesweep_kernel<2, ITEMS>;
    case 3: return radix_onesweep_kernel<3, ITEMS>;
    case 4: return radix_onesweep_kernel<4, ITEMS>;
    case 5: return radix_onesweep_kernel<5, ITEMS>;
    case 6: return radix_onesweep_kernel<6, ITEMS>;
    case 7: return radix_onesweep_kernel<7, ITEMS>;
    default: return radix_onesweep_kernel<8, ITEMS>;
  }
}

// Stable sort of `n` (u32 key, u32 value) pairs by key bits [0, end_bit). Input in
// keys[0]/values[0] (values[0] may be implied = element index when iota_values);
// *result_buf = index (0/1) of the buffer pair holding the result.
// `ghist` (passes*256 u32) must hold the digit COUNTS when have_hist, else zeros (it keeps the counts:
// the passes scan them themselves);
// `status` must be zero on entry.
int launch_radix_sort(uint32_t* keys[2], uint32_t* values[2], uint32_t n, int end_bit,
                      bool descending, bool iota_values, bool have_hist, uint32_t* ghist,
                      uint32_t* status, int n_sort_tiles, int* result_buf, cudaStream_t stream) {
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
  const uint32_t flip = descending ? ~0u : 0u;
  const int tiles = sort_tiles_for(n);
  const size_t smem = sort_smem_bytes(SORT_ITEMS);
  if (tiles > n_sort_tiles) {
    set_error("radix sort: status region sized for %d tiles, need %d", n_sort_tiles, tiles);
    return SOAR_E_WORKSPACE;
  }
  if (!have_hist) {
    const int hist_blocks = min(tiles * 2, 148 * 8);
    KernelScope ks(K_RADIX_HIST, stream);
    radix_hist_kernel<<<hist_blocks, 256, 0, stream>>>(keys[0], n, passes, end_bit, flip, ghist);
  }
  SOAR_LAUNCH_CHECK("radix_hist_kernel");
  uint32_t* tickets = status + (size_t)passes * n_sort_tiles * SORT_RADIX;
  // per device and idempotent (a host-side attribute write, no device work)
  for (int p = 0; p < passes; ++p) {
    const int shift = sort_pass_shift(p, end_bit, passes);
    const int bits = sort_pass_bits(p, end_bit, passes);
    OnesweepFn fn = onesweep_for_bits<SORT_ITEMS>(bits);
    // per device and idempotent (host-side attribute writes, no device work); 4 CTAs/SM (64
    // registers under the launch bound) need 4 x 35 KB of shared memory
    SOAR_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SOAR_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       cudaSharedmemCarveoutMaxShared));
    {
      KernelScope ks(K_RADIX_ONESWEEP, stream);
      fn<<<tiles, SORT_THREADS, smem, stream>>>(
          keys[p & 1], keys[(p + 1) & 1], (p == 0 && iota_values) ? nullptr : values[p & 1],
          values[(p + 1) & 1], n, shift, flip, ghist + p * SORT_RADIX,
          status + (size_t)p * n_sort_tiles * SORT_RADIX, tickets + p);
    }
    SOAR_LAUNCH_CHECK("radix_onesweep_kernel");
  }
  *result_buf = passes & 1;
  return SOAR_OK;
}

// =================================================================================
// Tile ranges (rasterizer_impl.cu:104-124) from the sorted tile ids
// =================================================================================
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t L, const uint32_t* __restrict__ keys, uint2* ranges, uint32_t n_tiles) {
  // four consecutive list entries per thread (one 16-byte load) + the entry before them
  const uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
  if (base >= L) return;
  uint32_t k[4];
  if (base + 4 <= L) {
    const uint4 v = *reinterpret_cast<const uint4*>(keys + base);
    k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) k[i] = base + i < L ? keys[base + i] : 0u;
  }
  uint32_t prev = base > 0 ? keys[base - 1] : 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t idx = base + i;
    if (idx < L) {
      const uint32_t cur = k[i];
      // ranges[n_tiles].x counts the non-empty tiles: the blend launches pick their layout by it
      if (idx == 0) {
        ranges[cur].x = 0;
        atomicAdd(&ranges[n_tiles].x, 1u);
      } else if (cur != prev) {
        ranges[prev].y = idx;
        ranges[cur].x = idx;
        atomicAdd(&ranges[n_tiles].x, 1u);
      }
      if (idx == L - 1) ranges[cur].y = L;
      prev = cur;
    }
  }
}

int launch_tile_ranges(uint32_t L, const uint32_t* keys, uint2* ranges, int n_tiles,
                       cudaStream_t stream) {
  SOAR_CUDA_TRY(cudaMemsetAsync(ranges, 0, ((size_t)n_tiles + 2) * sizeof(uint2), stream));
  if (L == 0) return SOAR_OK;
  {
    KernelScope ks(K_TILE_RANGES, stream);
    tile_ranges_kernel<<<ceil_div(ceil_div((int)L, 4), 256), 256, 0, stream>>>(L, keys, ranges, (uint32_t)n_tiles);
  }
  SOAR_LAUNCH_CHECK("tile_ranges_kernel");
  return SOAR_OK;
}

// =================================================================================
// Debug: materialise the reference's 64-bit keys (tile << 32 | depth bits) of the
// sorted list, for the bit-exact parity tests.
// =================================================================================
__global__ void __launch_bounds__(256)
debug_keys_kernel(uint32_t L, const uint32_t* __restrict__ tiles, const uint32_t* __restrict__ ids,
                  const float* __restrict__ records, unsigned long long* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const float depth = records[(size_t)ids[i] * REC_FLOATS + REC_DEPTH];
  out[i] = ((unsigned long long)tiles[i] << 32) | __float_as_uint(depth);
}

int launch_debug_keys(uint32_t L, const uint32_t* tiles, const uint32_t* ids, const float* records,
                      unsigned long long* out, cudaStream_t stream) {
  if (L == 0) return SOAR_OK;
  debug_keys_kernel<<<ceil_div((int)L, 256), 256, 0, stream>>>(L, tiles, ids, records, out);
  SOAR_LAUNCH_CHECK("debug_keys_kernel");
  return SOAR_OK;
}

}  // namespace soar
