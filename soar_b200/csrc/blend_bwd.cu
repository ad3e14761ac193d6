// K7 — tile alpha-blend backward for sm_100a.
//
// Behavioural contract: DGR/cuda_rasterizer/backward.cu:529-858. Per pixel the
// blend is replayed back-to-front from final_T / n_contrib; per contribution
// the same 13 partial derivatives as the reference are formed (colour 3,
// normal 3 (x10, Q4), depth 1, mean2D 2, conic 3, opacity 1).
//
// What is ours:
//   * the reference issues 13 global atomicAdds per (pixel, Gaussian) contribution. Here
//     the partials are reduced across the 32 pixels of a warp with a transposed butterfly
//     (12 or 16 shuffles for all 10 / 14 values together, not 5 per value) and the warp adds
//     its totals with ONE warp-wide reduction instruction into the Gaussian's 64-byte
//     accumulator line (10 / 14 consecutive words); the mean2D / conic gradients are formed
//     from moment sums once per Gaussian in geom_bwd.cu, not per contribution;
//   * the replay starts at the tile's deepest n_contrib instead of the end of its range;
//   * list entries that cannot reach alpha >= 1/255 anywhere in the tile are dropped at
//     staging time, and every warp (an 8x4 pixel block) only visits the entries whose
//     footprint can touch its block (blend_common.cuh); the block masks are the bytes the
//     forward blend left in the binning buffer, so a culled entry costs a one-byte load;
//   * pixel-constant factors are hoisted, 1/(1-alpha) is formed once per contribution, and
//     the body is specialised on (surface, per-pixel depth, normalize_depth).
#include <stdlib.h>
#include "common.cuh"
#include "layout.h"
#include "kernels.h"
#include "blend_common.cuh"

namespace soar {

constexpr int BWD_BATCH = TILE_PIX;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// The replay must take the forward's (= the reference's) per-pixel decision `alpha < 1/255` exactly: a
// contribution that flips in or out is a 0.4 % step in every gradient it feeds, and that — not the
// 2^-22 value error of ex2.approx — was the whole of round 1's 2e-4 deviation on opaque surfels
// (tests/test_grad_attribution_gpu.py). ex2.approx decides wrongly only inside a
// relative band of ~2^-21 around the threshold; opacity * G values inside a 16x wider band are
// recomputed with the forward's precise expf. One contribution in ~10^5 takes that path.
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float ALPHA_BAND = ALPHA_MIN * 8e-6f;
__device__ __forceinline__ bool near_alpha_threshold(float og) { return fabsf(og - ALPHA_MIN) < ALPHA_BAND; }

__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Sum v[0..15] over the 32 lanes; on return lane l holds the total of value
// index (l >> 1) in v[0] (both lanes of a pair hold the same total).
__device__ __forceinline__ float warp_reduce16(float (&v)[16], uint32_t lane) {
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float send = hi ? v[i] : v[i + 8];
      const float keep = hi ? v[i + 8] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool hi = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float send = hi ? v[i] : v[i + 4];
      const float keep = hi ? v[i + 4] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
  }
  {
    const bool hi = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float send = hi ? v[i] : v[i + 2];
      const float keep = hi ? v[i + 2] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
  }
  {
    const bool hi = lane & 2;
    const float send = hi ? v[0] : v[1];
    const float keep = hi ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
  return v[0];
}

// Slot order of the per-(pixel, Gaussian) partials = the accumulator's (layout.h, GaccSlot). The
// six dL/ddist moments replace the reference's five mean2D/conic sums: they are cheaper to form
// per pixel and need no per-tile conversion. Variants without surface normals only carry the
// first 10 slots (12-shuffle reduction instead of 16).
enum LocalSlot {
  LS_DC0 = GA_DC0, LS_DC1 = GA_DC1, LS_DC2 = GA_DC2, LS_DD = GA_DD, LS_DO = GA_DO, LS_MX = GA_MX,
  LS_MY = GA_MY, LS_MXX = GA_MXX, LS_MXY = GA_MXY, LS_MYY = GA_MYY, LS_DN0 = GA_DN0,
  LS_DN1 = GA_DN1, LS_DN2 = GA_DN2, LS_GD = GA_GD, LS_COUNT = GA_COUNT
};

// Sum v[0..9] over the 32 lanes with 12 shuffles. Returns the total of slot `idx` (output);
// idx < 0 on lanes that hold a duplicate.
__device__ __forceinline__ float warp_reduce10(float (&v)[16], uint32_t lane, int& idx) {
  const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4, h1 = lane & 2;
  // xor 16: pairs (i, i+5)
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const float send = h4 ? v[i] : v[i + 5];
    const float keep = h4 ? v[i + 5] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  // xor 8: pairs (0,3) (1,4); 2 is summed in place
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float send = h3 ? v[i] : v[i + 3];
    const float keep = h3 ? v[i + 3] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  v[2] += __shfl_xor_sync(0xffffffffu, v[2], 8);
  // xor 4: pair (0,1); 2 in place
  {
    const float send = h2 ? v[0] : v[1];
    const float keep = h2 ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
  v[2] += __shfl_xor_sync(0xffffffffu, v[2], 4);
  // xor 2: pair (0,2)
  {
    const float send = h1 ? v[0] : v[2];
    const float keep = h1 ? v[2] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
  const int base = h4 ? 5 : 0;
  if (!h1) idx = base + (h3 ? 3 : 0) + (h2 ? 1 : 0);
  else idx = (!h3 && !h2) ? base + 2 : -1;
  if (lane & 1) idx = -1;
  return v[0];
}

struct BlendBwdSmem {
  float4 q0[BWD_BATCH];  // x y conA conB
  float4 q1[BWD_BATCH];  // conC opac thr depth
  float4 q2[BWD_BATCH];  // r g b J6
  float4 q3[BWD_BATCH];  // nx ny nz J9
  float4 q4[BWD_BATCH];  // J0 J1 J2 J3
  uint32_t id[BWD_BATCH];
  uint32_t pos[BWD_BATCH];  // 0-based position in the tile's list
  uint8_t mask[BWD_BATCH];
  uint32_t wcount[N_BLK];
  uint32_t max_contrib[N_BLK];
};

// MODE selects an arithmetic variant of the scalar twin. The variants exist to ATTRIBUTE the
// gradient deviation from the reference (tests/test_grad_attribution_gpu.py):
//   bit 0 (1)  precise expf and IEEE division instead of ex2.approx / rcp.approx
//   bit 1 (2)  the reference's per-channel accum_rec recurrences and separate T_final terms instead
//              of the single scalar S recurrence and the fused tf_k constant
//   bit 2 (4)  dL/dmean2D summed per pixel with the reference's expression
//              dd * (A dx + B dy), dd * (C dy + B dx) instead of being formed at flush time from the
//              moments sum(dd dx), sum(dd dy) (BlendBwdArgs.mean_direct tells geom_bwd which it got)
template <bool SURFACE, bool DEPTH_DIFF, bool NORM_DEPTH, int MODE>
__device__ __forceinline__ void blend_bwd_body(const BlendBwdArgs& a, BlendBwdSmem& sm, uint32_t tile_x,
                                               uint32_t tile_y) {
  constexpr bool PRECISE = MODE & 1, PER_CHANNEL = MODE & 2, MEAN_DIRECT = MODE & 4;
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  const uint32_t px = tile_x * TILE_X + pix_x_in_tile(warp, lane);
  const uint32_t py = tile_y * TILE_Y + pix_y_in_tile(warp, lane);
  const bool inside = px < (uint32_t)a.W && py < (uint32_t)a.H;
  const uint32_t pix_id = (uint32_t)a.W * py + px;
  const float2 pixf = {(float)px, (float)py};
  const size_t HW = (size_t)a.H * a.W;

  const uint2 range = a.ranges[tile_y * a.grid_x + tile_x];
  if (range.y == range.x) return;

  // lanes outside the image never contribute; 0.5 keeps their (discarded) arithmetic finite
  const float T_final = inside ? a.final_T[pix_id] : 0.5f;
  const float D_final = inside && NORM_DEPTH ? a.final_D[pix_id] : 0;
  const uint32_t last_contributor = inside ? a.n_contrib[pix_id] : 0;

  // deepest contributor of the tile (and of this warp's block): nothing behind it receives
  // a gradient
  uint32_t block_depth = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    block_depth = max(block_depth, __shfl_xor_sync(0xffffffffu, block_depth, o));
  if (lane == 0) sm.max_contrib[warp] = block_depth;
  __syncthreads();
  uint32_t tile_depth = 0;
#pragma unroll
  for (int w = 0; w < N_BLK; ++w) tile_depth = max(tile_depth, sm.max_contrib[w]);
  if (tile_depth == 0) return;
  // the tile's list, restricted to [range.x, range.x + tile_depth), walked from the back
  const uint32_t list_end = range.x + tile_depth;
  const int rounds = (tile_depth + BWD_BATCH - 1) / BWD_BATCH;

  float dL_dpixC[3] = {0.f, 0.f, 0.f}, dL_dpixN[3] = {0.f, 0.f, 0.f}, dL_dpixD = 0.f, dL_dpixO = 0.f;
  if (inside) {
#pragma unroll
    for (int i = 0; i < 3; i++) dL_dpixC[i] = a.dL_dpixcolor[i * HW + pix_id];
    if (SURFACE) {
#pragma unroll
      for (int i = 0; i < 3; i++) dL_dpixN[i] = a.dL_dpixnormal[i * HW + pix_id];
    }
    dL_dpixD = a.dL_dpixdepth[pix_id] * 1;
    dL_dpixO = a.dL_dpixopac[pix_id];
  }
  float bg_dot_dpixel = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) bg_dot_dpixel += a.bg[i] * dL_dpixC[i];

  // pixel constants of the depth terms (backward.cu:760-777): the channel gradient is divided
  // by (1 - T_final) when depth is normalised, and the normalisation adds
  // g_d * D_final / (1-T_f)^2 * (-T_f) / ((1-alpha) * T) to dL/dalpha.
  float dL_ddepth_ch = dL_dpixD, norm_k = 0.f;
  if (NORM_DEPTH) {
    dL_ddepth_ch = dL_dpixD / (1.f - T_final);
    norm_k = dL_dpixD * D_final / (1.f - T_final) / (1.f - T_final) * -T_final;
  }
  // background / opacity / un-normalised depth terms all scale with T_final / (1 - alpha)
  float tf_k = dL_dpixO * T_final - T_final * bg_dot_dpixel;
  if (!NORM_DEPTH) tf_k -= T_final * (10 * dL_dpixD);

  float T = T_final;
  float S = 0.f, cg_last = 0.f, last_alpha = 0.f;
  float accum_rec[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, last_ch[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};

  for (int i = 0; i < rounds; i++) {
    // ---- stage (from the back), cull against the tile, compact in walk order ---------
    const uint32_t progress = i * BWD_BATCH + tid;
    uint32_t mask = 0, id = 0;
    const float4* rec = nullptr;
    float4 q0, q1;
    if (progress < tile_depth) {
      // every entry in front of the tile's deepest contributor was staged by the forward,
      // which left its block mask behind: culled entries cost one byte here
      mask = a.pair_mask[list_end - progress - 1];
      if (mask != 0) {
        id = a.point_list[list_end - progress - 1];
        rec = reinterpret_cast<const float4*>(a.records + (size_t)id * REC_FLOATS);
        q0 = __ldg(rec + 0);
        q1 = __ldg(rec + 1);
      }
    }
    const uint32_t keep_ballot = __ballot_sync(0xffffffffu, mask != 0);
    if (lane == 0) sm.wcount[warp] = __popc(keep_ballot);
    __syncthreads();
    uint32_t slot = __popc(keep_ballot & lanemask_lt()), n = 0;
#pragma unroll
    for (uint32_t w = 0; w < N_BLK; ++w) {
      const uint32_t c = sm.wcount[w];
      if (w < warp) slot += c;
      n += c;
    }
    if (mask != 0) {
      sm.q0[slot] = q0;
      sm.q1[slot] = q1;
      sm.q2[slot] = __ldg(rec + 2);
      if (SURFACE) sm.q3[slot] = __ldg(rec + 3);
      if (DEPTH_DIFF) sm.q4[slot] = __ldg(rec + 4);
      sm.id[slot] = id;
      sm.pos[slot] = tile_depth - 1 - progress;
      sm.mask[slot] = (uint8_t)mask;
    }
    __syncthreads();

    for (uint32_t c0 = 0; c0 < n; c0 += 32) {
      const uint32_t jj = c0 + lane;
      const bool mine = jj < n && ((sm.mask[jj] >> warp) & 1u) && sm.pos[jj] < block_depth;
      uint32_t bits = __ballot_sync(0xffffffffu, mine);
      while (bits) {
        const uint32_t j = c0 + __ffs(bits) - 1;
        bits &= bits - 1;
        bool live = sm.pos[j] < last_contributor;
        const float4 q0j = sm.q0[j];
        const float4 q1j = sm.q1[j];
        const float2 d = {q0j.x - pixf.x, q0j.y - pixf.y};
        const float dist = (q0j.z * d.x * d.x + q1j.x * d.y * d.y) + 2 * q0j.w * d.x * d.y;
        const float power = -0.5f * dist;
        live = live && !(power > 0.0f) && !(power < q1j.z);
        if (!__any_sync(0xffffffffu, live)) continue;
        // the backward is compared at 1e-3 relative (the reference's own atomics are unordered):
        // single-instruction ex2 / rcp approximations (2^-22 relative) are far inside that
        float G = PRECISE ? expf(power) : ex2_approx(power * 1.4426950408889634f);
        if (!PRECISE && live && near_alpha_threshold(q1j.y * G)) G = expf(power);
        float alpha = min(0.99f, q1j.y * G);
        live = live && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, live)) continue;

        // Lanes that do not contribute run the same arithmetic with alpha = 0 and G = 0: T, S and
        // every partial then come out unchanged / zero, so nothing below is predicated.
        alpha = live ? alpha : 0.f;
        const float Gl = live ? G : 0.f;
        float v[16];
#pragma unroll
        for (int k = LS_COUNT; k < 16; ++k) v[k] = 0.f;
        {
          const float T_behind = T;  // transmittance after this Gaussian (front-to-back)
          const float rcp = PRECISE ? 1.f / (1.f - alpha) : rcp_approx(1.f - alpha);
          T = PRECISE ? T / (1.f - alpha) : T * rcp;
          const float dchannel_dcolor = alpha * T;
          const float4 q2 = sm.q2[j];

          // dL/dalpha needs sum_ch g_ch * (c_ch - accum_rec_ch) with
          // accum_rec_ch <- last_alpha * last_c_ch + (1 - last_alpha) * accum_rec_ch for the 3
          // colour, 3 normal and 1 depth channels (backward.cu:700-770). The recurrence is
          // linear, so S = sum_ch g_ch * accum_rec_ch is carried as ONE scalar:
          // S <- last_alpha * cg_last + (1 - last_alpha) * S, cg = sum_ch g_ch * c_ch.
          float cg = q2.x * dL_dpixC[0] + q2.y * dL_dpixC[1] + q2.z * dL_dpixC[2];
          v[LS_DC0] = dchannel_dcolor * dL_dpixC[0];
          v[LS_DC1] = dchannel_dcolor * dL_dpixC[1];
          v[LS_DC2] = dchannel_dcolor * dL_dpixC[2];
          float d_cur = q1j.w;
          v[LS_DN0] = v[LS_DN1] = v[LS_DN2] = v[LS_GD] = 0.f;
          if (SURFACE) {
            const float4 q3 = sm.q3[j];
            cg += q3.x * dL_dpixN[0] + q3.y * dL_dpixN[1] + q3.z * dL_dpixN[2];
            v[LS_DN0] = dchannel_dcolor * dL_dpixN[0] * 10;
            v[LS_DN1] = dchannel_dcolor * dL_dpixN[1] * 10;
            v[LS_DN2] = dchannel_dcolor * dL_dpixN[2] * 10;
            if (DEPTH_DIFF) {
              const float4 q4 = sm.q4[j];
              const float dif_u0 = d.x * q4.x + d.y * q4.y;
              const float dif_u1 = d.x * q4.z + d.y * q4.w;
              d_cur -= dif_u0 * q2.w + dif_u1 * q3.w;
              v[LS_GD] = live ? dL_dpixD : 0.f;  // the pixel-depth term of dL/dmean2D is formed at flush
            }
          }
          cg += d_cur * dL_ddepth_ch;
          v[LS_DD] = dchannel_dcolor * dL_ddepth_ch * 1;
          float dL_dalpha;
          if (!PER_CHANNEL) {
            S = last_alpha * cg_last + (1.f - last_alpha) * S;
            cg_last = cg;
            dL_dalpha = cg - S;
            if (NORM_DEPTH) dL_dalpha += norm_k / T_behind;  // = norm_k / ((1-alpha) * T)
            dL_dalpha *= T;
            dL_dalpha += tf_k * rcp;
          } else if (live) {
            // backward.cu:700-790 term by term (skipped contributions leave every recurrence
            // untouched there, hence the predicate)
            const float4 q3 = SURFACE ? sm.q3[j] : make_float4(0.f, 0.f, 0.f, 0.f);
            const float c_cur[7] = {q2.x, q2.y, q2.z, q3.x, q3.y, q3.z, d_cur};
            const float g_ch[7] = {dL_dpixC[0], dL_dpixC[1], dL_dpixC[2], dL_dpixN[0], dL_dpixN[1],
                                   dL_dpixN[2], dL_ddepth_ch};
            dL_dalpha = 0.f;
#pragma unroll
            for (int ch = 0; ch < 7; ++ch) {
              if (!SURFACE && ch >= 3 && ch < 6) continue;
              accum_rec[ch] = last_alpha * last_ch[ch] + (1.f - last_alpha) * accum_rec[ch];
              last_ch[ch] = c_cur[ch];
              float part = 0.f;
              if (ch == 6 && NORM_DEPTH)
                part += dL_dpixD * D_final / (1.f - T_final) / (1.f - T_final) * -T_final / (1 - alpha) / T;
              part += (c_cur[ch] - accum_rec[ch]) * g_ch[ch];
              dL_dalpha += part;
            }
            dL_dalpha *= T;
            dL_dalpha += dL_dpixO * T_final / (1 - alpha);
            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
            if (!NORM_DEPTH) dL_dalpha += (-T_final / (1.f - alpha)) * (10 * dL_dpixD);
          } else {
            dL_dalpha = 0.f;
          }
          if (!PER_CHANNEL || live) last_alpha = alpha;

          const float gda = Gl * dL_dalpha;
          const float dL_ddist = gda * q1j.y * -0.5f;
          const float ddx = dL_ddist * d.x, ddy = dL_ddist * d.y;
          v[LS_DO] = gda;
          if (MEAN_DIRECT) {
            v[LS_MX] = dL_ddist * (q0j.z * d.x + q0j.w * d.y);
            v[LS_MY] = dL_ddist * (q1j.x * d.y + q0j.w * d.x);
          } else {
            v[LS_MX] = ddx;
            v[LS_MY] = ddy;
          }
          v[LS_MXX] = ddx * d.x;
          v[LS_MXY] = ddx * d.y;
          v[LS_MYY] = ddy * d.y;
        }

        // one warp-wide RED: the lanes holding a total hit consecutive words of the Gaussian's
        // 64-byte accumulator line (native fp32 reduction at L2; shared-memory float atomics are
        // compare-and-swap loops)
        float* const g = a.gacc + (size_t)sm.id[j] * GACC_FLOATS;
        if (SURFACE) {
          const float total = warp_reduce16(v, lane);
          if ((lane & 1) == 0 && (lane >> 1) < LS_COUNT) atomicAdd(g + (lane >> 1), total);
        } else {
          int slot_idx;
          const float total = warp_reduce10(v, lane, slot_idx);
          if (slot_idx >= 0) atomicAdd(g + slot_idx, total);
        }
      }
    }
    // no barrier here: the next round only overwrites the batch after its own first barrier
  }
}

// ---------------------------------------------------------------------------------------------
// Two pixels per thread (the forward's blend_fwd2 layout: 128-thread CTAs, warp = 8x8 pixel block =
// two 8x4 mask blocks, lane (lx, ly) replays pixels (lx, ly) and (lx, ly + 4)) with packed f32x2
// arithmetic. Per (Gaussian, 8x8 block) the loop control, the shared loads and — the point of it —
// the warp reduction of the 10 / 14 partials are paid ONCE for 64 pixels instead of once per 32: the
// two pixels' partials are added in the lane before the butterfly. Same formulas as blend_bwd_body;
// the sums are associated differently (pixel pair first), which is inside the 1e-3 contract.
// ---------------------------------------------------------------------------------------------
constexpr int BWD2_THREADS = TILE_PIX / 2;

__device__ __forceinline__ float2 bc(float s) { return make_float2(s, s); }

template <bool SURFACE, bool DEPTH_DIFF, bool NORM_DEPTH>
__device__ __forceinline__ void blend_bwd2_body(const BlendBwdArgs& a, BlendBwdSmem& sm) {
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_x = blockIdx.x, tile_y = blockIdx.y;
  const uint32_t px = tile_x * TILE_X + (warp & 1u) * 8u + (lane & 7u);
  const uint32_t py0 = tile_y * TILE_Y + (warp >> 1) * 8u + (lane >> 3);
  const uint32_t py1 = py0 + 4u;
  const bool inside0 = px < (uint32_t)a.W && py0 < (uint32_t)a.H;
  const bool inside1 = px < (uint32_t)a.W && py1 < (uint32_t)a.H;
  const uint32_t pix0 = (uint32_t)a.W * py0 + px, pix1 = (uint32_t)a.W * py1 + px;
  const float pxf = (float)px;
  const float2 npyf = make_float2(-(float)py0, -(float)py1);
  const size_t HW = (size_t)a.H * a.W;
  const uint32_t wA = 4u * (warp >> 1) + (warp & 1u);
  const uint32_t my_blocks = 5u << wA;

  const uint2 range = a.ranges[tile_y * a.grid_x + tile_x];
  if (range.y == range.x) return;

  // lanes outside the image never contribute; 0.5 keeps their (discarded) arithmetic finite
  const float2 T_final = make_float2(inside0 ? a.final_T[pix0] : 0.5f, inside1 ? a.final_T[pix1] : 0.5f);
  const float2 D_final = make_float2(inside0 && NORM_DEPTH ? a.final_D[pix0] : 0.f,
                                     inside1 && NORM_DEPTH ? a.final_D[pix1] : 0.f);
  const uint32_t last0 = inside0 ? a.n_contrib[pix0] : 0, last1 = inside1 ? a.n_contrib[pix1] : 0;

  uint32_t block_depth = max(last0, last1);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    block_depth = max(block_depth, __shfl_xor_sync(0xffffffffu, block_depth, o));
  if (lane == 0) sm.max_contrib[warp] = block_depth;
  __syncthreads();
  uint32_t tile_depth = 0;
#pragma unroll
  for (int w = 0; w < BWD2_THREADS / 32; ++w) tile_depth = max(tile_depth, sm.max_contrib[w]);
  if (tile_depth == 0) return;
  const uint32_t list_end = range.x + tile_depth;
  const int rounds = (tile_depth + BWD_BATCH - 1) / BWD_BATCH;

  float2 gC0 = bc(0.f), gC1 = gC0, gC2 = gC0, gN0 = gC0, gN1 = gC0, gN2 = gC0, gD = gC0, gO = gC0;
  if (inside0) {
    gC0.x = a.dL_dpixcolor[0 * HW + pix0]; gC1.x = a.dL_dpixcolor[1 * HW + pix0]; gC2.x = a.dL_dpixcolor[2 * HW + pix0];
    if (SURFACE) { gN0.x = a.dL_dpixnormal[0 * HW + pix0]; gN1.x = a.dL_dpixnormal[1 * HW + pix0]; gN2.x = a.dL_dpixnormal[2 * HW + pix0]; }
    gD.x = a.dL_dpixdepth[pix0] * 1;
    gO.x = a.dL_dpixopac[pix0];
  }
  if (inside1) {
    gC0.y = a.dL_dpixcolor[0 * HW + pix1]; gC1.y = a.dL_dpixcolor[1 * HW + pix1]; gC2.y = a.dL_dpixcolor[2 * HW + pix1];
    if (SURFACE) { gN0.y = a.dL_dpixnormal[0 * HW + pix1]; gN1.y = a.dL_dpixnormal[1 * HW + pix1]; gN2.y = a.dL_dpixnormal[2 * HW + pix1]; }
    gD.y = a.dL_dpixdepth[pix1] * 1;
    gO.y = a.dL_dpixopac[pix1];
  }
  float2 bg_dot = bc(0.f);
  bg_dot = __ffma2_rn(bc(a.bg[0]), gC0, bg_dot);
  bg_dot = __ffma2_rn(bc(a.bg[1]), gC1, bg_dot);
  bg_dot = __ffma2_rn(bc(a.bg[2]), gC2, bg_dot);

  // pixel constants of the depth terms, as in blend_bwd_body
  float2 g_depth_ch = gD, norm_k = bc(0.f);
  if (NORM_DEPTH) {
    g_depth_ch = make_float2(gD.x / (1.f - T_final.x), gD.y / (1.f - T_final.y));
    norm_k = make_float2(gD.x * D_final.x / (1.f - T_final.x) / (1.f - T_final.x) * -T_final.x,
                         gD.y * D_final.y / (1.f - T_final.y) / (1.f - T_final.y) * -T_final.y);
  }
  float2 tf_k = make_float2(gO.x * T_final.x - T_final.x * bg_dot.x, gO.y * T_final.y - T_final.y * bg_dot.y);
  if (!NORM_DEPTH) {
    tf_k.x -= T_final.x * (10 * gD.x);
    tf_k.y -= T_final.y * (10 * gD.y);
  }

  float2 T = T_final, S = bc(0.f), cg_last = bc(0.f), last_alpha = bc(0.f);

  for (int i = 0; i < rounds; i++) {
    // ---- stage two entries per thread (from the back), compact in walk order ---------------
    uint32_t mask[2] = {0, 0}, id[2] = {0, 0};
    const float4* rec[2] = {nullptr, nullptr};
    float4 q0[2], q1[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const uint32_t progress = i * BWD_BATCH + h * BWD2_THREADS + tid;
      if (progress < tile_depth) {
        mask[h] = a.pair_mask[list_end - progress - 1];
        if (mask[h] != 0) {
          id[h] = a.point_list[list_end - progress - 1];
          rec[h] = reinterpret_cast<const float4*>(a.records + (size_t)id[h] * REC_FLOATS);
          q0[h] = __ldg(rec[h] + 0);
          q1[h] = __ldg(rec[h] + 1);
        }
      }
    }
    uint32_t keep_ballot[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      keep_ballot[h] = __ballot_sync(0xffffffffu, mask[h] != 0);
      if (lane == 0) sm.wcount[h * 4 + warp] = __popc(keep_ballot[h]);
    }
    __syncthreads();
    uint32_t slot[2] = {__popc(keep_ballot[0] & lanemask_lt()), __popc(keep_ballot[1] & lanemask_lt())};
    uint32_t n = 0;
#pragma unroll
    for (uint32_t w = 0; w < N_BLK; ++w) {
      const uint32_t c = sm.wcount[w];
      if (w < warp) slot[0] += c;
      if (w < 4u + warp) slot[1] += c;
      n += c;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (mask[h] != 0) {
        const uint32_t sl = slot[h];
        sm.q0[sl] = q0[h];
        sm.q1[sl] = q1[h];
        sm.q2[sl] = __ldg(rec[h] + 2);
        if (SURFACE) sm.q3[sl] = __ldg(rec[h] + 3);
        if (DEPTH_DIFF) sm.q4[sl] = __ldg(rec[h] + 4);
        sm.id[sl] = id[h];
        sm.pos[sl] = tile_depth - 1 - (i * BWD_BATCH + h * BWD2_THREADS + tid);
        sm.mask[sl] = (uint8_t)mask[h];
      }
    }
    __syncthreads();

    for (uint32_t c0 = 0; c0 < n; c0 += 32) {
      const uint32_t jj = c0 + lane;
      const bool mine = jj < n && (sm.mask[jj] & my_blocks) && sm.pos[jj] < block_depth;
      uint32_t bits = __ballot_sync(0xffffffffu, mine);
      while (bits) {
        const uint32_t j = c0 + __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t pos = sm.pos[j];
        bool live0 = pos < last0, live1 = pos < last1;
        const float4 q0j = sm.q0[j];
        const float4 q1j = sm.q1[j];
        const float dx = q0j.x - pxf;
        const float2 dy = __fadd2_rn(bc(q0j.y), npyf);
        const float adx = q0j.z * dx;
        const float bdx = dx * (q0j.w + q0j.w);
        const float2 cdy = __fmul2_rn(bc(q1j.x), dy);
        float2 dist = __fmul2_rn(dy, cdy);
        dist = __ffma2_rn(bc(dx), bc(adx), dist);
        dist = __ffma2_rn(dy, bc(bdx), dist);
        const float2 power = __fmul2_rn(dist, bc(-0.5f));
        live0 = live0 && !(power.x > 0.0f) && !(power.x < q1j.z);
        live1 = live1 && !(power.y > 0.0f) && !(power.y < q1j.z);
        if (!__any_sync(0xffffffffu, live0 || live1)) continue;
        const float2 pl = __fmul2_rn(power, bc(1.4426950408889634f));
        float2 G = make_float2(ex2_approx(pl.x), ex2_approx(pl.y));
        float2 og = __fmul2_rn(bc(q1j.y), G);
        {
          const bool n0 = live0 && near_alpha_threshold(og.x), n1 = live1 && near_alpha_threshold(og.y);
          if (__any_sync(0xffffffffu, n0 || n1)) {   // ~1e-5 of the contributions: decide exactly
            if (n0) { G.x = expf(power.x); og.x = q1j.y * G.x; }
            if (n1) { G.y = expf(power.y); og.y = q1j.y * G.y; }
          }
        }
        float2 alpha = make_float2(min(0.99f, og.x), min(0.99f, og.y));
        live0 = live0 && !(alpha.x < 1.0f / 255.0f);
        live1 = live1 && !(alpha.y < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, live0 || live1)) continue;

        // lanes that do not contribute run the same arithmetic with alpha = 0 and G = 0
        alpha.x = live0 ? alpha.x : 0.f;
        alpha.y = live1 ? alpha.y : 0.f;
        const float2 Gl = make_float2(live0 ? G.x : 0.f, live1 ? G.y : 0.f);
        float v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = 0.f;
        {
          const float2 T_behind = T;
          const float2 oma = __fadd2_rn(bc(1.f), make_float2(-alpha.x, -alpha.y));
          const float2 rcp = make_float2(rcp_approx(oma.x), rcp_approx(oma.y));
          T = __fmul2_rn(T, rcp);
          const float2 dch = __fmul2_rn(alpha, T);
          const float4 q2 = sm.q2[j];

          float2 cg = __fmul2_rn(bc(q2.x), gC0);
          cg = __ffma2_rn(bc(q2.y), gC1, cg);
          cg = __ffma2_rn(bc(q2.z), gC2, cg);
          const float2 vC0 = __fmul2_rn(dch, gC0), vC1 = __fmul2_rn(dch, gC1), vC2 = __fmul2_rn(dch, gC2);
          v[LS_DC0] = vC0.x + vC0.y;
          v[LS_DC1] = vC1.x + vC1.y;
          v[LS_DC2] = vC2.x + vC2.y;
          float2 d_cur = bc(q1j.w);
          if (SURFACE) {
            const float4 q3 = sm.q3[j];
            cg = __ffma2_rn(bc(q3.x), gN0, cg);
            cg = __ffma2_rn(bc(q3.y), gN1, cg);
            cg = __ffma2_rn(bc(q3.z), gN2, cg);
            const float2 dch10 = __fmul2_rn(dch, bc(10.f));
            const float2 vN0 = __fmul2_rn(dch10, gN0), vN1 = __fmul2_rn(dch10, gN1), vN2 = __fmul2_rn(dch10, gN2);
            v[LS_DN0] = vN0.x + vN0.y;
            v[LS_DN1] = vN1.x + vN1.y;
            v[LS_DN2] = vN2.x + vN2.y;
            if (DEPTH_DIFF) {
              const float4 q4 = sm.q4[j];
              // dif_u0 = dx J0 + dy J1, dif_u1 = dx J2 + dy J3, d_cur -= dif_u0 J6 + dif_u1 J9
              const float2 u0 = __ffma2_rn(dy, bc(q4.y), bc(dx * q4.x));
              const float2 u1 = __ffma2_rn(dy, bc(q4.w), bc(dx * q4.z));
              const float2 dz = __ffma2_rn(u1, bc(q3.w), __fmul2_rn(u0, bc(q2.w)));
              d_cur = __fadd2_rn(d_cur, make_float2(-dz.x, -dz.y));
              v[LS_GD] = (live0 ? gD.x : 0.f) + (live1 ? gD.y : 0.f);
            }
          }
          cg = __ffma2_rn(d_cur, g_depth_ch, cg);
          const float2 vDD = __fmul2_rn(dch, g_depth_ch);
          v[LS_DD] = vDD.x + vDD.y;
          // S <- last_alpha * cg_last + (1 - last_alpha) * S
          const float2 oml = __fadd2_rn(bc(1.f), make_float2(-last_alpha.x, -last_alpha.y));
          S = __ffma2_rn(last_alpha, cg_last, __fmul2_rn(oml, S));
          cg_last = cg;
          float2 dL_dalpha = __fadd2_rn(cg, make_float2(-S.x, -S.y));
          if (NORM_DEPTH) {
            dL_dalpha.x += norm_k.x / T_behind.x;
            dL_dalpha.y += norm_k.y / T_behind.y;
          }
          dL_dalpha = __fmul2_rn(dL_dalpha, T);
          dL_dalpha = __ffma2_rn(tf_k, rcp, dL_dalpha);
          last_alpha = alpha;

          const float2 gda = __fmul2_rn(Gl, dL_dalpha);
          const float2 dL_ddist = __fmul2_rn(__fmul2_rn(gda, bc(q1j.y)), bc(-0.5f));
          const float2 ddx = __fmul2_rn(dL_ddist, bc(dx)), ddy = __fmul2_rn(dL_ddist, dy);
          const float2 mxx = __fmul2_rn(ddx, bc(dx)), mxy = __fmul2_rn(ddx, dy), myy = __fmul2_rn(ddy, dy);
          v[LS_DO] = gda.x + gda.y;
          v[LS_MX] = ddx.x + ddx.y;
          v[LS_MY] = ddy.x + ddy.y;
          v[LS_MXX] = mxx.x + mxx.y;
          v[LS_MXY] = mxy.x + mxy.y;
          v[LS_MYY] = myy.x + myy.y;
        }

        float* const g = a.gacc + (size_t)sm.id[j] * GACC_FLOATS;
        if (SURFACE) {
          const float total = warp_reduce16(v, lane);
          if ((lane & 1) == 0 && (lane >> 1) < LS_COUNT) atomicAdd(g + (lane >> 1), total);
        } else {
          int slot_idx;
          const float total = warp_reduce10(v, lane, slot_idx);
          if (slot_idx >= 0) atomicAdd(g + slot_idx, total);
        }
      }
    }
  }
}

// The product kernel: one instantiation per (surface, per-pixel depth, normalize_depth) combination the
// reference's config can take, picked on the HOST when the caller knows the flags
// (SoarRastView.config_host_flags) — each body then gets the whole 64-register budget to itself. The
// all-in-one kernel below, which branches on the device copy of config (backward.cu:581), serves callers
// that only hold the device tensor; it spills (the allocator has to cover all six bodies at once).
template <bool SURFACE, bool DEPTH_DIFF, bool NORM_DEPTH>
// 7 CTAs/SM (73 registers) measured 343.6 us at c3 against 358.7 at 8 (64 registers) and 350.4 at 6
// (tools/gpu_bounds_sweep.sh): the ninth register row buys more than the eighth CTA.
__global__ void __launch_bounds__(BWD2_THREADS, 7) blend_bwd2_kernel_t(const BlendBwdArgs a) {
  blend_exit_if_layout_declines(a.ranges, a.n_tiles, a.sparse_below, false);
  __shared__ BlendBwdSmem sm;
  blend_bwd2_body<SURFACE, DEPTH_DIFF, NORM_DEPTH>(a, sm);
}

template <bool SURFACE, bool DEPTH_DIFF>
__device__ __forceinline__ void blend_bwd2_pick(const BlendBwdArgs& a, BlendBwdSmem& sm, bool normalize_depth) {
  if (normalize_depth) blend_bwd2_body<SURFACE, DEPTH_DIFF, true>(a, sm);
  else blend_bwd2_body<SURFACE, DEPTH_DIFF, false>(a, sm);
}

__global__ void __launch_bounds__(BWD2_THREADS, 8) blend_bwd2_kernel(const BlendBwdArgs a) {
  blend_exit_if_layout_declines(a.ranges, a.n_tiles, a.sparse_below, false);
  __shared__ BlendBwdSmem sm;
  const bool surface = a.config[0] > 0, per_pixel_depth = a.config[2] > 0,
             normalize_depth = a.config[1] > 0;
  if (surface && per_pixel_depth) blend_bwd2_pick<true, true>(a, sm, normalize_depth);
  else if (surface) blend_bwd2_pick<true, false>(a, sm, normalize_depth);
  else blend_bwd2_pick<false, false>(a, sm, normalize_depth);
}

// The scalar twin (one pixel per thread), kept for the twin tests and the attribution variants.
template <int MODE>
__global__ void __launch_bounds__(TILE_PIX, 4) blend_bwd_kernel(const BlendBwdArgs a) {
  blend_exit_if_layout_declines(a.ranges, a.n_tiles, a.sparse_below, true);
  __shared__ BlendBwdSmem sm;
  const bool surface = a.config[0] > 0, per_pixel_depth = a.config[2] > 0,
             normalize_depth = a.config[1] > 0;
  if (surface && per_pixel_depth) {
    if (normalize_depth) blend_bwd_body<true, true, true, MODE>(a, sm, blockIdx.x, blockIdx.y);
    else blend_bwd_body<true, true, false, MODE>(a, sm, blockIdx.x, blockIdx.y);
  } else if (surface) {
    if (normalize_depth) blend_bwd_body<true, false, true, MODE>(a, sm, blockIdx.x, blockIdx.y);
    else blend_bwd_body<true, false, false, MODE>(a, sm, blockIdx.x, blockIdx.y);
  } else {
    if (normalize_depth) blend_bwd_body<false, false, true, MODE>(a, sm, blockIdx.x, blockIdx.y);
    else blend_bwd_body<false, false, false, MODE>(a, sm, blockIdx.x, blockIdx.y);
  }
}

// One pixel per thread, MODE 0, one instantiation per config combination picked on the host: the kernel
// small launches take (kernels.h, blend_prefers_one_pixel_threads).
template <bool SURFACE, bool DEPTH_DIFF, bool NORM_DEPTH>
__global__ void __launch_bounds__(TILE_PIX, 4) blend_bwd1_kernel_t(const BlendBwdArgs a) {
  blend_exit_if_layout_declines(a.ranges, a.n_tiles, a.sparse_below, true);
  __shared__ BlendBwdSmem sm;
  blend_bwd_body<SURFACE, DEPTH_DIFF, NORM_DEPTH, 0>(a, sm, blockIdx.x, blockIdx.y);
}

// Kernel choice of a launch (kernels.h): environment override, else by the size of the launch.
int blend_prefers_one_pixel_threads(const char* var, uint32_t num_rendered) {
  const char* e = getenv(var);
  if (e) {
    const int n = atoi(e);
    if (n <= 0) return -1;
    return n <= 8 ? n - 1 : 0;
  }
  return num_rendered <= BLEND_SMALL_LAUNCH_PAIRS ? 0 : -1;
}
// Non-empty tiles at or below which a frame counts as sparse: the packed kernels' resident CTA slots
// (8 per SM) filled less than twice. 0 = an environment override decides instead.
uint32_t blend_sparse_threshold(const char* var) {
  if (getenv(var)) return 0;
  const char* t = getenv("SOAR_BLEND_SPARSE_TILES");   // experiments
  if (t) return (uint32_t)atoi(t);
  static int sms = 0;
  if (sms == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    sms = n;
  }
  return (uint32_t)sms * 8u * 2u;
}
int blend_bwd_scalar_mode(uint32_t num_rendered) {
  return blend_prefers_one_pixel_threads("SOAR_BLEND_BWD_SCALAR", num_rendered);
}

// (see blend_fwd1p_kernel_t) The counter is the .x of the second statistics slot behind the ranges.
template <bool SURFACE, bool DEPTH_DIFF, bool NORM_DEPTH>
__global__ void __launch_bounds__(TILE_PIX, 4) blend_bwd1p_kernel_t(const BlendBwdArgs a) {
  blend_exit_if_layout_declines(a.ranges, a.n_tiles, a.sparse_below, true);
  __shared__ BlendBwdSmem sm;
  __shared__ uint32_t s_next;
  uint32_t* const counter = const_cast<uint32_t*>(&a.ranges[a.n_tiles + 1].x);
  for (;;) {
    if (threadIdx.x == 0) s_next = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t t = s_next;
    __syncthreads();
    if (t >= a.n_tiles) break;
    blend_bwd_body<SURFACE, DEPTH_DIFF, NORM_DEPTH, 0>(a, sm, t % (uint32_t)a.grid_x, t / (uint32_t)a.grid_x);
  }
}
static void launch_bwd_one_pixel_drawn(const BlendBwdArgs& a, int ctas, cudaStream_t stream) {
  const bool surface = a.cfg_flags & SOAR_CFG_SURFACE, norm = a.cfg_flags & SOAR_CFG_NORMALIZE_DEPTH,
             dd = surface && (a.cfg_flags & SOAR_CFG_PERPIX_DEPTH);
#define SOAR_BWD1P(S, D, N) blend_bwd1p_kernel_t<S, D, N><<<ctas, TILE_PIX, 0, stream>>>(a)
  if (dd) { if (norm) SOAR_BWD1P(true, true, true); else SOAR_BWD1P(true, true, false); }
  else if (surface) { if (norm) SOAR_BWD1P(true, false, true); else SOAR_BWD1P(true, false, false); }
  else { if (norm) SOAR_BWD1P(false, false, true); else SOAR_BWD1P(false, false, false); }
#undef SOAR_BWD1P
}

static void launch_bwd_one_pixel(const BlendBwdArgs& a, dim3 grid, int mode, cudaStream_t stream) {
  if (mode == 0 && (a.cfg_flags & SOAR_CFG_KNOWN)) {
    const bool surface = a.cfg_flags & SOAR_CFG_SURFACE, norm = a.cfg_flags & SOAR_CFG_NORMALIZE_DEPTH,
               dd = surface && (a.cfg_flags & SOAR_CFG_PERPIX_DEPTH);
#define SOAR_BWD1(S, D, N) blend_bwd1_kernel_t<S, D, N><<<grid, TILE_PIX, 0, stream>>>(a)
    if (dd) { if (norm) SOAR_BWD1(true, true, true); else SOAR_BWD1(true, true, false); }
    else if (surface) { if (norm) SOAR_BWD1(true, false, true); else SOAR_BWD1(true, false, false); }
    else { if (norm) SOAR_BWD1(false, false, true); else SOAR_BWD1(false, false, false); }
#undef SOAR_BWD1
    return;
  }
  switch (mode) {
    case 0: blend_bwd_kernel<0><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 1: blend_bwd_kernel<1><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 2: blend_bwd_kernel<2><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 3: blend_bwd_kernel<3><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 4: blend_bwd_kernel<4><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 5: blend_bwd_kernel<5><<<grid, TILE_PIX, 0, stream>>>(a); break;
    case 6: blend_bwd_kernel<6><<<grid, TILE_PIX, 0, stream>>>(a); break;
    default: blend_bwd_kernel<7><<<grid, TILE_PIX, 0, stream>>>(a); break;
  }
}
static void launch_bwd_packed(const BlendBwdArgs& a, dim3 grid, cudaStream_t stream) {
  if (a.cfg_flags & SOAR_CFG_KNOWN) {
    const bool surface = a.cfg_flags & SOAR_CFG_SURFACE, norm = a.cfg_flags & SOAR_CFG_NORMALIZE_DEPTH,
               dd = surface && (a.cfg_flags & SOAR_CFG_PERPIX_DEPTH);
#define SOAR_BWD2(S, D, N) blend_bwd2_kernel_t<S, D, N><<<grid, BWD2_THREADS, 0, stream>>>(a)
    if (dd) { if (norm) SOAR_BWD2(true, true, true); else SOAR_BWD2(true, true, false); }
    else if (surface) { if (norm) SOAR_BWD2(true, false, true); else SOAR_BWD2(true, false, false); }
    else { if (norm) SOAR_BWD2(false, false, true); else SOAR_BWD2(false, false, false); }
#undef SOAR_BWD2
  } else {
    blend_bwd2_kernel<<<grid, BWD2_THREADS, 0, stream>>>(a);
  }
}

int launch_blend_bwd(const BlendBwdArgs& a_in, int grid_y, cudaStream_t stream) {
  BlendBwdArgs a = a_in;
  dim3 grid(a.grid_x, grid_y, 1);
  {
    KernelScope ks(K_BLEND_BWD, stream);
    const int mode = blend_bwd_scalar_mode(a.num_rendered);
    const uint32_t sparse_below = (a.cfg_flags & SOAR_CFG_KNOWN) ? blend_sparse_threshold("SOAR_BLEND_BWD_SCALAR") : 0;
    a.sparse_below = 0;
    if (mode >= 0) {
      launch_bwd_one_pixel(a, grid, mode, stream);   // small launch, or forced
    } else if (sparse_below == 0) {
      launch_bwd_packed(a, grid, stream);            // forced
    } else {
      a.sparse_below = sparse_below;                 // the frame decides (kernels.h)
      // the tile counter of the drawn launch: a backward may run more than once on one forward
      if (cudaMemsetAsync(const_cast<uint32_t*>(&a.ranges[a.n_tiles + 1].x), 0, sizeof(uint32_t), stream) != cudaSuccess) {
        set_error("blend_bwd: resetting the tile counter failed");
        return SOAR_E_CUDA;
      }
      launch_bwd_packed(a, grid, stream);
      KernelScope::count_extra_launch();
      launch_bwd_one_pixel_drawn(a, (int)max(1u, min(a.n_tiles, sparse_below / 4u)), stream);
    }
  }
  SOAR_LAUNCH_CHECK("blend_bwd_kernel");
  return SOAR_OK;
}

}  // namespace soar
