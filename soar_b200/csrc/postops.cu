// Image-space post-ops of SOAR's surfel renderer as ONE stencil kernel forward and ONE backward
// (SURVEY.md §8 row f2).
//
// Behavioural contract (torch op chains in the reference, ~25 launches + a CPU torch.inverse
// with a blocking .to(device) per render):
//   * depth2normal(depth, mask, camera)      TS/renderer/diff_gaussian_rasterizer.py:359-430
//       camPos = (((w - cx)·D)·Kinv00, ((h - cy)·D)·Kinv11, D), replicate-padded 5-point stencil,
//       n = u×l + r×u + b×r + l×b with every arm masked, F.normalize(eps 1e-12), × mask.
//       Quirk kept: K00 is built from FoVy/height and scales x, K11 from FoVx/width scales y (:378-380).
//   * normal2curv(normal, mask)              :433-448
//       s = Σ arms (n[nb] - n·m_c)·m_nb, × mask, L1 norm over the channels.
//   * the renderer's post block              :292-303
//       mask = opac > 1e-5; normal gradient cut outside the mask; channels 1,2 of the normal and
//       of the depth normal negated; (x + 1)/2 on both; curvature of the flipped normal.
//
// One CTA = a 32×8 pixel tile; depth / flipped normal / mask are staged once with a halo
// (1 forward, 2 backward) using clamped coordinates, which IS the replicate padding.
// The backward is a gather (no atomics, deterministic): a pixel q collects from the five stencils
// it takes part in — its own and those of its four neighbours, plus its own again for every arm
// that the padding folds back onto q at an image border.
// HBM-bound: 49 B/pixel forward, 61 B/pixel backward.
#include "common.cuh"

namespace soar {

constexpr int PT_X = 32;
constexpr int PT_Y = 8;
constexpr int PT_THREADS = PT_X * PT_Y;

struct PostCtx {
  int W, H;
  float cx, cy, kx, ky, thr;
  int surfel;
  const uint8_t* mask;
  const float* opac;
  const float* normal;
  const float* depth;
};

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator-(const float3& a, const float3& b) {
  return f3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z));
}
__device__ __forceinline__ float3 operator+(const float3& a, const float3& b) {
  return f3(a.x + b.x, a.y + b.y, a.z + b.z);
}
__device__ __forceinline__ float3 sel(bool m, const float3& a) { return m ? a : f3(0.f, 0.f, 0.f); }
__device__ __forceinline__ float3 cross3(const float3& a, const float3& b) {
  return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

template <int HALO>
struct Tile {
  static constexpr int SW = PT_X + 2 * HALO;
  static constexpr int SH = PT_Y + 2 * HALO;
  static constexpr int N = SW * SH;
};

// Stages mask / depth / (flipped) normal of the tile + halo; coordinates are clamped to the image,
// so a halo cell outside it holds its replicate-padded value.
template <int HALO>
__device__ __forceinline__ void stage(const PostCtx& c, int x0, int y0, uint8_t* sM, float* sD,
                                      float* sN) {
  using T = Tile<HALO>;
  const size_t npix = (size_t)c.W * c.H;
  const float sg = c.surfel ? -1.f : 1.f;
  for (int i = threadIdx.x; i < T::N; i += PT_THREADS) {
    const int lx = i % T::SW, ly = i / T::SW;
    const int gx = min(max(x0 + lx - HALO, 0), c.W - 1);
    const int gy = min(max(y0 + ly - HALO, 0), c.H - 1);
    const size_t g = (size_t)gy * c.W + gx;
    sM[i] = c.mask ? (c.mask[g] != 0) : (__ldg(c.opac + g) > c.thr);
    if (c.depth) sD[i] = __ldg(c.depth + g);
    if (c.normal) {
      sN[i] = __ldg(c.normal + g);
      sN[T::N + i] = sg * __ldg(c.normal + npix + g);
      sN[2 * T::N + i] = sg * __ldg(c.normal + 2 * npix + g);
    }
  }
}

// Explicitly rounded products: a folded arm (border replication) must be EXACTLY zero, as it is in
// the reference; an fma contracted across `P_nb - P_c` would leave a rounding residue that
// F.normalize then blows up to a unit vector on degenerate stencils.
__device__ __forceinline__ float3 cam_pos(const PostCtx& c, int gx, int gy, float D) {
  return f3(__fmul_rn(__fmul_rn((float)gx - c.cx, D), c.kx),
            __fmul_rn(__fmul_rn((float)gy - c.cy, D), c.ky), D);
}

struct Arms {
  float3 u, l, b, r;
  bool mc, mu, ml, mb, mr;
};

// The four masked arms of depth2normal's stencil at image pixel (gx, gy) = staged cell `i`.
template <int SW>
__device__ __forceinline__ Arms depth_arms(const PostCtx& c, const uint8_t* sM, const float* sD,
                                           int i, int gx, int gy) {
  Arms a;
  a.mc = sM[i]; a.mu = sM[i - SW]; a.ml = sM[i - 1]; a.mb = sM[i + SW]; a.mr = sM[i + 1];
  const int gxl = max(gx - 1, 0), gxr = min(gx + 1, c.W - 1);
  const int gyu = max(gy - 1, 0), gyb = min(gy + 1, c.H - 1);
  const float3 pc = sel(a.mc, cam_pos(c, gx, gy, sD[i]));
  a.u = sel(a.mu, cam_pos(c, gx, gyu, sD[i - SW]) - pc);
  a.l = sel(a.ml, cam_pos(c, gxl, gy, sD[i - 1]) - pc);
  a.b = sel(a.mb, cam_pos(c, gx, gyb, sD[i + SW]) - pc);
  a.r = sel(a.mr, cam_pos(c, gxr, gy, sD[i + 1]) - pc);
  return a;
}

__device__ __forceinline__ float3 arms_normal(const Arms& a) {
  // n_ul + n_ur + n_br + n_bl, in the reference's order (:407-420)
  return ((cross3(a.u, a.l) + cross3(a.r, a.u)) + cross3(a.b, a.r)) + cross3(a.l, a.b);
}

template <int SW, int N>
__device__ __forceinline__ float3 ld3(const float* s, int i) {
  return f3(s[i], s[N + i], s[2 * N + i]);
}

// normal2curv's summed arms at staged cell i, before the final mask / L1 norm.
template <int SW, int N>
__device__ __forceinline__ float3 curv_sum(const uint8_t* sM, const float* sN, int i) {
  const float3 nc = sel(sM[i], ld3<SW, N>(sN, i));
  const float3 nu = sel(sM[i - SW], ld3<SW, N>(sN, i - SW) - nc);
  const float3 nl = sel(sM[i - 1], ld3<SW, N>(sN, i - 1) - nc);
  const float3 nb = sel(sM[i + SW], ld3<SW, N>(sN, i + SW) - nc);
  const float3 nr = sel(sM[i + 1], ld3<SW, N>(sN, i + 1) - nc);
  return ((nu + nl) + nb) + nr;
}

__global__ void __launch_bounds__(PT_THREADS)
post_fwd_kernel(PostCtx c, float* __restrict__ out_normal, float* __restrict__ out_curv,
                float* __restrict__ out_dnormal, uint8_t* __restrict__ out_mask) {
  using T = Tile<1>;
  __shared__ uint8_t sM[T::N];
  __shared__ float sD[T::N];
  __shared__ float sN[3 * T::N];
  const int x0 = blockIdx.x * PT_X, y0 = blockIdx.y * PT_Y;
  stage<1>(c, x0, y0, sM, sD, sN);
  __syncthreads();
  const int tx = threadIdx.x % PT_X, ty = threadIdx.x / PT_X;
  const int gx = x0 + tx, gy = y0 + ty;
  if (gx >= c.W || gy >= c.H) return;
  const int i = (ty + 1) * T::SW + tx + 1;
  const size_t npix = (size_t)c.W * c.H, g = (size_t)gy * c.W + gx;
  const bool mc = sM[i];
  if (out_mask) out_mask[g] = mc;
  if (c.normal) {
    if (out_curv) {
      const float3 s = sel(mc, curv_sum<T::SW, T::N>(sM, sN, i));
      out_curv[g] = (fabsf(s.x) + fabsf(s.y)) + fabsf(s.z);
    }
    if (out_normal) {  // surfel block: (flipped normal + 1) / 2
      const float3 n = ld3<T::SW, T::N>(sN, i);
      out_normal[g] = (n.x + 1.f) * 0.5f;
      out_normal[npix + g] = (n.y + 1.f) * 0.5f;
      out_normal[2 * npix + g] = (n.z + 1.f) * 0.5f;
    }
  }
  if (c.depth && out_dnormal) {
    const Arms a = depth_arms<T::SW>(c, sM, sD, i, gx, gy);
    float3 n = arms_normal(a);
    const float len = sqrtf((n.x * n.x + n.y * n.y) + n.z * n.z);
    const float den = fmaxf(len, 1e-12f);
    n = sel(mc, f3(n.x / den, n.y / den, n.z / den));
    if (c.surfel) {
      n = f3((n.x + 1.f) * 0.5f, (-n.y + 1.f) * 0.5f, (-n.z + 1.f) * 0.5f);
    }
    out_dnormal[g] = n.x;
    out_dnormal[npix + g] = n.y;
    out_dnormal[2 * npix + g] = n.z;
  }
}

// dL/d(arm) of depth2normal for the four arms, given G = dL/dn (unnormalised n).
//   n = u×l + r×u + b×r + l×b  =>  dL/du = (l - r)×G, dL/dl = (b - u)×G, dL/dr = (u - b)×G,
//   dL/db = (r - l)×G
struct ArmGrads {
  float3 u, l, b, r;
};
__device__ __forceinline__ ArmGrads arm_grads(const Arms& a, const float3& G) {
  ArmGrads d;
  d.u = cross3(a.l - a.r, G);
  d.l = cross3(a.b - a.u, G);
  d.r = cross3(a.u - a.b, G);
  d.b = cross3(a.r - a.l, G);
  return d;
}

__global__ void __launch_bounds__(PT_THREADS)
post_bwd_kernel(PostCtx c, const float* __restrict__ g_out_normal, const float* __restrict__ g_curv,
                const float* __restrict__ g_dnormal, float* __restrict__ g_normal,
                float* __restrict__ g_depth) {
  using T2 = Tile<2>;  // inputs
  using T1 = Tile<1>;  // per-stencil cotangents
  __shared__ uint8_t sM[T2::N];
  __shared__ float sD[T2::N];
  __shared__ float sN[3 * T2::N];
  __shared__ float sG[3 * T1::N];  // dL/dn (depth2normal, unnormalised)
  __shared__ float sS[3 * T1::N];  // dL/ds (normal2curv)
  const int x0 = blockIdx.x * PT_X, y0 = blockIdx.y * PT_Y;
  const size_t npix = (size_t)c.W * c.H;
  stage<2>(c, x0, y0, sM, sD, sN);
  __syncthreads();
  const bool do_d = c.depth && g_dnormal && g_depth;
  const bool do_c = c.normal && g_curv && g_normal;
  // stage 1: the cotangent of every stencil centred in tile + halo 1 (zero outside the image)
  for (int j = threadIdx.x; j < T1::N; j += PT_THREADS) {
    const int lx = j % T1::SW, ly = j / T1::SW;
    const int gx = x0 + lx - 1, gy = y0 + ly - 1;
    const bool inside = gx >= 0 && gx < c.W && gy >= 0 && gy < c.H;
    const int i = (ly + 1) * T2::SW + lx + 1;
    float3 G = f3(0.f, 0.f, 0.f), S = f3(0.f, 0.f, 0.f);
    if (inside) {
      const size_t g = (size_t)gy * c.W + gx;
      const bool mc = sM[i];
      if (do_d && mc) {
        float3 go = f3(__ldg(g_dnormal + g), __ldg(g_dnormal + npix + g),
                       __ldg(g_dnormal + 2 * npix + g));
        if (c.surfel) go = f3(0.5f * go.x, -0.5f * go.y, -0.5f * go.z);
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i, gx, gy);
        const float3 n = arms_normal(a);
        const float len = sqrtf((n.x * n.x + n.y * n.y) + n.z * n.z);
        if (len >= 1e-12f) {
          const float inv = 1.f / len;
          const float3 nh = f3(n.x * inv, n.y * inv, n.z * inv);
          const float dt = nh.x * go.x + nh.y * go.y + nh.z * go.z;
          G = f3((go.x - nh.x * dt) * inv, (go.y - nh.y * dt) * inv, (go.z - nh.z * dt) * inv);
        } else {
          G = f3(go.x * 1e12f, go.y * 1e12f, go.z * 1e12f);
        }
      }
      if (do_c && mc) {
        const float gc = __ldg(g_curv + g);
        const float3 s = curv_sum<T2::SW, T2::N>(sM, sN, i);
        S = f3(s.x > 0.f ? gc : (s.x < 0.f ? -gc : 0.f), s.y > 0.f ? gc : (s.y < 0.f ? -gc : 0.f),
               s.z > 0.f ? gc : (s.z < 0.f ? -gc : 0.f));
      }
    }
    sG[j] = G.x; sG[T1::N + j] = G.y; sG[2 * T1::N + j] = G.z;
    sS[j] = S.x; sS[T1::N + j] = S.y; sS[2 * T1::N + j] = S.z;
  }
  __syncthreads();
  // stage 2: gather
  const int tx = threadIdx.x % PT_X, ty = threadIdx.x / PT_X;
  const int gx = x0 + tx, gy = y0 + ty;
  if (gx >= c.W || gy >= c.H) return;
  const int i = (ty + 2) * T2::SW + tx + 2;  // in the halo-2 arrays
  const int j = (ty + 1) * T1::SW + tx + 1;  // in the halo-1 arrays
  const size_t g = (size_t)gy * c.W + gx;
  const bool mq = sM[i];
  const bool top = gy == 0, left = gx == 0, bottom = gy == c.H - 1, right = gx == c.W - 1;

  if (do_d) {
    float3 acc = f3(0.f, 0.f, 0.f);
    if (mq) {
      {  // q's own stencil: centre role, and arms folded back onto q by the padding
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i, gx, gy);
        const ArmGrads d = arm_grads(a, ld3<T1::SW, T1::N>(sG, j));
        const float3 du = sel(a.mu, d.u), dl = sel(a.ml, d.l), db = sel(a.mb, d.b),
                     dr = sel(a.mr, d.r);
        acc = acc - (((du + dl) + db) + dr);
        if (top) acc = acc + du;
        if (left) acc = acc + dl;
        if (bottom) acc = acc + db;
        if (right) acc = acc + dr;
      }
      if (!bottom) {  // q is the upper neighbour of the pixel below
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i + T2::SW, gx, gy + 1);
        acc = acc + cross3(a.l - a.r, ld3<T1::SW, T1::N>(sG, j + T1::SW));
      }
      if (!right) {  // q is the left neighbour of the pixel to the right
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i + 1, gx + 1, gy);
        acc = acc + cross3(a.b - a.u, ld3<T1::SW, T1::N>(sG, j + 1));
      }
      if (!top) {  // q is the lower neighbour of the pixel above
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i - T2::SW, gx, gy - 1);
        acc = acc + cross3(a.r - a.l, ld3<T1::SW, T1::N>(sG, j - T1::SW));
      }
      if (!left) {  // q is the right neighbour of the pixel to the left
        const Arms a = depth_arms<T2::SW>(c, sM, sD, i - 1, gx - 1, gy);
        acc = acc + cross3(a.u - a.b, ld3<T1::SW, T1::N>(sG, j - 1));
      }
    }
    g_depth[g] = (acc.x * (((float)gx - c.cx) * c.kx) + acc.y * (((float)gy - c.cy) * c.ky)) + acc.z;
  }

  if (c.normal && g_normal) {
    float3 acc = f3(0.f, 0.f, 0.f);
    if (do_c && mq) {
      const float3 Sq = ld3<T1::SW, T1::N>(sS, j);
      const int cnt = (int)sM[i - T2::SW] + (int)sM[i - 1] + (int)sM[i + T2::SW] + (int)sM[i + 1];
      const int fold = (int)top + (int)left + (int)bottom + (int)right;  // those arms see mask m_q = 1
      const float k = (float)(fold - cnt);
      acc = f3(k * Sq.x, k * Sq.y, k * Sq.z);
      if (!bottom) acc = acc + ld3<T1::SW, T1::N>(sS, j + T1::SW);
      if (!right) acc = acc + ld3<T1::SW, T1::N>(sS, j + 1);
      if (!top) acc = acc + ld3<T1::SW, T1::N>(sS, j - T1::SW);
      if (!left) acc = acc + ld3<T1::SW, T1::N>(sS, j - 1);
    }
    if (c.surfel) {
      // out_normal = (n' + 1)/2, n' = (n0, -n1, -n2); gradient cut outside the mask (:292-294)
      if (g_out_normal) {
        acc.x += 0.5f * __ldg(g_out_normal + g);
        acc.y += 0.5f * __ldg(g_out_normal + npix + g);
        acc.z += 0.5f * __ldg(g_out_normal + 2 * npix + g);
      }
      acc = sel(mq, f3(acc.x, -acc.y, -acc.z));
    }
    g_normal[g] = acc.x;
    g_normal[npix + g] = acc.y;
    g_normal[2 * npix + g] = acc.z;
  }
}

static int make_ctx(const SoarPostArgs* a, PostCtx* c, const char* who) {
  if (!a) { set_error("%s: args is NULL", who); return SOAR_E_INVALID; }
  if (a->width <= 0 || a->height <= 0) {
    set_error("%s: bad image size %dx%d", who, a->width, a->height);
    return SOAR_E_INVALID;
  }
  if (!a->mask && !a->opac) {
    set_error("%s: needs a mask or an opacity image", who);
    return SOAR_E_INVALID;
  }
  if (!a->normal && !a->depth) {
    set_error("%s: needs a normal or a depth image", who);
    return SOAR_E_INVALID;
  }
  c->W = a->width; c->H = a->height;
  c->cx = a->cx; c->cy = a->cy; c->kx = a->kinv_x; c->ky = a->kinv_y;
  c->thr = a->mask_threshold; c->surfel = a->surfel;
  c->mask = a->mask; c->opac = a->opac; c->normal = a->normal; c->depth = a->depth;
  return SOAR_OK;
}

}  // namespace soar

using namespace soar;

extern "C" {

int soar_post_forward(const SoarPostArgs* args, float* out_normal, float* out_curv,
                      float* out_depth_normal, uint8_t* out_mask, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PostCtx c;
  if (int rc = make_ctx(args, &c, "soar_post_forward")) return rc;
  if (out_normal && !c.surfel) {
    set_error("soar_post_forward: out_normal exists only in the surfel block (surfel=1)");
    return SOAR_E_INVALID;
  }
  dim3 grid(ceil_div(c.W, PT_X), ceil_div(c.H, PT_Y));
  {
    KernelScope ks(K_POST_FWD, stream);
    post_fwd_kernel<<<grid, PT_THREADS, 0, stream>>>(c, out_normal, out_curv, out_depth_normal,
                                                     out_mask);
  }
  SOAR_LAUNCH_CHECK("post_fwd_kernel");
  return SOAR_OK;
}

int soar_post_backward(const SoarPostArgs* args, const float* dL_dout_normal, const float* dL_dcurv,
                       const float* dL_ddepth_normal, float* dL_dnormal, float* dL_ddepth,
                       soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PostCtx c;
  if (int rc = make_ctx(args, &c, "soar_post_backward")) return rc;
  if (dL_ddepth && (!c.depth || !dL_ddepth_normal)) {
    set_error("soar_post_backward: dL_ddepth needs depth and dL_ddepth_normal");
    return SOAR_E_INVALID;
  }
  if (dL_dnormal && (!c.normal || !(dL_dcurv || (c.surfel && dL_dout_normal)))) {
    set_error("soar_post_backward: dL_dnormal needs normal and an incoming gradient");
    return SOAR_E_INVALID;
  }
  dim3 grid(ceil_div(c.W, PT_X), ceil_div(c.H, PT_Y));
  {
    KernelScope ks(K_POST_BWD, stream);
    post_bwd_kernel<<<grid, PT_THREADS, 0, stream>>>(c, dL_dout_normal, dL_dcurv, dL_ddepth_normal,
                                                     dL_dnormal, dL_ddepth);
  }
  SOAR_LAUNCH_CHECK("post_bwd_kernel");
  return SOAR_OK;
}

}  // extern "C"
