// K7 — tile alpha-blend backward for sm_100a.
//
// Behavioural contract: DGR/cuda_rasterizer/backward.cu:529-858. Per pixel the
// blend is replayed back-to-front from final_T / n_contrib; per contribution
// the same 13 partial derivatives as the reference are formed (colour 3,
// normal 3 (x10, Q4), depth 1, mean2D 2, conic 3, opacity 1).
//
// What is ours: the reference issues 13 global atomicAdds per (pixel, Gaussian)
// contribution. Here the 13 values are reduced across the 32 pixels of a warp
// with a transposed butterfly (16 shuffles for all 13 values together, not 5
// per value), the 8 warps of the tile meet in a shared-memory accumulator, and
// the tile flushes ONE set of (vectorised) global reductions per Gaussian per
// batch into a 64-byte per-Gaussian accumulator. Warps in which no pixel sees
// the Gaussian skip the reduction with one vote; the replay starts at the
// tile's deepest n_contrib instead of the end of the tile's range.
#include "common.cuh"
#include "layout.h"
#include "kernels.h"

namespace soar {

constexpr int BWD_BATCH = TILE_PIX;


__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
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

__global__ void __launch_bounds__(TILE_PIX) blend_bwd_kernel(const BlendBwdArgs a) {
  __shared__ float4 s_q0[BWD_BATCH];  // x y conA conB
  __shared__ float4 s_q1[BWD_BATCH];  // conC opac depth J6
  __shared__ float4 s_q2[BWD_BATCH];  // r g b J9
  __shared__ float4 s_q3[BWD_BATCH];  // nx ny nz -
  __shared__ float4 s_q4[BWD_BATCH];  // J0 J1 J2 J3
  __shared__ uint32_t s_id[BWD_BATCH];
  __shared__ float s_acc[BWD_BATCH][GACC_FLOATS];
  __shared__ uint32_t s_max_contrib[TILE_PIX / 32];

  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_x = blockIdx.x, tile_y = blockIdx.y;
  const uint32_t px = tile_x * TILE_X + (tid & (TILE_X - 1));
  const uint32_t py = tile_y * TILE_Y + (tid >> 4);
  const bool inside = px < (uint32_t)a.W && py < (uint32_t)a.H;
  const uint32_t pix_id = (uint32_t)a.W * py + px;
  const float2 pixf = {(float)px, (float)py};
  const size_t HW = (size_t)a.H * a.W;

  const uint2 range = a.ranges[tile_y * a.grid_x + tile_x];
  if (range.y == range.x) return;

  const bool surface = a.config[0] > 0, per_pixel_depth = a.config[2] > 0,
             normalize_depth = a.config[1] > 0;
  const bool depth_diff = surface && per_pixel_depth;

  const float T_final = inside ? a.final_T[pix_id] : 0;
  const float D_final = inside && normalize_depth ? a.final_D[pix_id] : 0;
  const uint32_t last_contributor = inside ? a.n_contrib[pix_id] : 0;

  // deepest contributor of the tile: nothing behind it receives a gradient
  {
    uint32_t m = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_max_contrib[warp] = m;
  }
#pragma unroll
  for (int k = 0; k < GACC_FLOATS; ++k) s_acc[tid][k] = 0.f;
  __syncthreads();
  uint32_t tile_depth = 0;
#pragma unroll
  for (int w = 0; w < TILE_PIX / 32; ++w) tile_depth = max(tile_depth, s_max_contrib[w]);
  if (tile_depth == 0) return;
  // the tile's list, restricted to [range.x, range.x + tile_depth), walked from the back
  const uint32_t list_end = range.x + tile_depth;
  int toDo = (int)tile_depth;
  const int rounds = (toDo + BWD_BATCH - 1) / BWD_BATCH;

  float dL_dpixC[3] = {0.f, 0.f, 0.f}, dL_dpixN[3] = {0.f, 0.f, 0.f}, dL_dpixD = 0.f, dL_dpixO = 0.f;
  if (inside) {
#pragma unroll
    for (int i = 0; i < 3; i++) dL_dpixC[i] = a.dL_dpixcolor[i * HW + pix_id];
#pragma unroll
    for (int i = 0; i < 3; i++) dL_dpixN[i] = a.dL_dpixnormal[i * HW + pix_id];
    dL_dpixD = a.dL_dpixdepth[pix_id] * 1;
    dL_dpixO = a.dL_dpixopac[pix_id];
  }
  float bg_dot_dpixel = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) bg_dot_dpixel += a.bg[i] * dL_dpixC[i];

  float T = T_final;
  uint32_t contributor = tile_depth;
  float accum_rec[3] = {0.f, 0.f, 0.f}, accum_rec_n[3] = {0.f, 0.f, 0.f}, accum_rec_d = 0.f;
  float last_alpha = 0.f;
  float last_color[3] = {0.f, 0.f, 0.f}, last_normal[3] = {0.f, 0.f, 0.f}, last_depth = 0.f;
  const float ddelx_dx = 0.5 * a.W;
  const float ddely_dy = 0.5 * a.H;

  for (int i = 0; i < rounds; i++, toDo -= BWD_BATCH) {
    const uint32_t progress = i * BWD_BATCH + tid;
    const bool have = progress < tile_depth;
    if (have) {
      const uint32_t id = a.point_list[list_end - progress - 1];
      const float4* rec = reinterpret_cast<const float4*>(a.records + (size_t)id * REC_FLOATS);
      s_id[tid] = id;
      s_q0[tid] = __ldg(rec + 0);
      s_q1[tid] = __ldg(rec + 1);
      s_q2[tid] = __ldg(rec + 2);
      s_q3[tid] = __ldg(rec + 3);
      s_q4[tid] = __ldg(rec + 4);
    }
    __syncthreads();

    const int n = min(BWD_BATCH, toDo);
    for (int j = 0; j < n; j++) {
      contributor--;
      bool live = contributor < last_contributor;
      // warp-uniform early out before touching shared memory
      if (!__any_sync(0xffffffffu, live)) continue;

      const float4 q0 = s_q0[j];
      const float4 q1 = s_q1[j];
      const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
      const float dist = (q0.z * d.x * d.x + q1.x * d.y * d.y) + 2 * q0.w * d.x * d.y;
      const float power = -0.5f * dist;
      live = live && !(power > 0.0f);
      const float G = exp(power);
      float alpha = min(0.99f, q1.y * G);
      live = live && !(alpha < 1.0f / 255.0f);
      if (!__any_sync(0xffffffffu, live)) continue;

      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = 0.f;

      if (live) {
        T = T / (1.f - alpha);
        const float dchannel_dcolor = alpha * T;
        const float4 q2 = s_q2[j];
        const float4 q4 = depth_diff ? s_q4[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float J6 = depth_diff ? q1.w : 0.f, J9 = depth_diff ? q2.w : 0.f;

        float dL_dalpha = 0.0f;
        const float col[3] = {q2.x, q2.y, q2.z};
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          const float c_cur = col[ch];
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = c_cur;
          const float dL_dchannel = dL_dpixC[ch];
          dL_dalpha += (c_cur - accum_rec[ch]) * dL_dchannel;
          v[GA_DC0 + ch] = dchannel_dcolor * dL_dchannel;
        }
        if (surface) {
          const float4 q3 = s_q3[j];
          const float nrm[3] = {q3.x, q3.y, q3.z};
#pragma unroll
          for (int ch = 0; ch < 3; ch++) {
            const float n_cur = nrm[ch];
            accum_rec_n[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_rec_n[ch];
            last_normal[ch] = n_cur;
            const float dL_dchannel = dL_dpixN[ch];
            dL_dalpha += (n_cur - accum_rec_n[ch]) * dL_dchannel;
            v[GA_DN0 + ch] = dchannel_dcolor * dL_dchannel * 10;
          }
        }
        {
          float d_cur = q1.z;
          if (depth_diff) {
            const float dif_u0 = d.x * q4.x + d.y * q4.y;
            const float dif_u1 = d.x * q4.z + d.y * q4.w;
            d_cur -= dif_u0 * J6 + dif_u1 * J9;
          }
          accum_rec_d = last_alpha * last_depth + (1.f - last_alpha) * accum_rec_d;
          last_depth = d_cur;
          float dL_dchannel = dL_dpixD, dL_dalpha_depth = 0;
          if (normalize_depth) {
            dL_dchannel /= (1.f - T_final);
            dL_dalpha_depth +=
                dL_dpixD * D_final / (1.f - T_final) / (1.f - T_final) * -T_final / (1 - alpha) / T;
          }
          dL_dalpha_depth += (d_cur - accum_rec_d) * dL_dchannel;
          v[GA_DD] = dchannel_dcolor * dL_dchannel * 1;
          dL_dalpha += dL_dalpha_depth;
        }
        dL_dalpha *= T;
        dL_dalpha += dL_dpixO * T_final / (1 - alpha);
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
        if (!normalize_depth) dL_dalpha += (-T_final / (1.f - alpha)) * (10 * dL_dpixD);

        const float dL_ddist = dL_dalpha * q1.y * -0.5f * G;
        float ndc_x = dL_ddist * 2 * (q0.z * d.x + q0.w * d.y) * ddelx_dx;
        float ndc_y = dL_ddist * 2 * (q1.x * d.y + q0.w * d.x) * ddely_dy;
        if (depth_diff) {
          ndc_x += 1 * -dL_dpixD * (J6 * q4.x + J9 * q4.z);
          ndc_y += 1 * -dL_dpixD * (J6 * q4.y + J9 * q4.w);
        }
        v[GA_DMX] = ndc_x;
        v[GA_DMY] = ndc_y;
        v[GA_DCA] = dL_ddist * (d.x * d.x);
        v[GA_DCB] = dL_ddist * (1 * d.x * d.y);
        v[GA_DCC] = dL_ddist * (d.y * d.y);
        v[GA_DO] = G * dL_dalpha;
      }

      const float total = warp_reduce16(v, lane);
      if ((lane & 1) == 0 && (lane >> 1) < 13) atomicAdd(&s_acc[j][lane >> 1], total);
    }
    __syncthreads();

    // flush this batch: thread t owns the Gaussian it staged
    if (have) {
      float* acc = s_acc[tid];
      bool any = false;
#pragma unroll
      for (int k = 0; k < 13; ++k) any = any || (acc[k] != 0.f);
      if (any) {
        float* g = a.gacc + (size_t)s_id[tid] * GACC_FLOATS;
        red_add_v4(g + 0, acc[0], acc[1], acc[2], acc[3]);
        red_add_v4(g + 4, acc[4], acc[5], acc[6], acc[7]);
        red_add_v4(g + 8, acc[8], acc[9], acc[10], acc[11]);
        atomicAdd(g + 12, acc[12]);
#pragma unroll
        for (int k = 0; k < 13; ++k) acc[k] = 0.f;
      }
    }
    // next iteration's staging writes s_q*/s_id only after this barrier
    __syncthreads();
  }
}

int launch_blend_bwd(const BlendBwdArgs& a, int grid_y, cudaStream_t stream) {
  dim3 grid(a.grid_x, grid_y, 1);
  {
    KernelScope ks(K_BLEND_BWD, stream);
    blend_bwd_kernel<<<grid, TILE_PIX, 0, stream>>>(a);
  }
  SOAR_LAUNCH_CHECK("blend_bwd_kernel");
  return SOAR_OK;
}

}  // namespace soar
