// Optimizer step and densify / clone / split / prune of the Gaussian parameter groups for sm_100a
// (SURVEY.md §8 row f3).
//
// Behavioural contract (python/torch in the reference, TS/geometry/surfel_base.py):
//   * torch.optim.Adam(l, lr=0.0, eps=1e-15) over the per-Gaussian groups           :596-681
//   * add_densification_stats + max_radii2D update, once per rendered view            :1113-1135, :1214
//   * adaptive_densify / densify_and_clone / densify_and_split selection rules        :1083-1111, :979-1067
//   * adaptive_prune rule                                                              :1069-1081
//   * prune_points / _prune_optimizer (boolean-mask row selection of every parameter,
//     both Adam moments and the six statistics)                                       :862-903
//   * cat_tensors_to_optimizer / densification_postfix (append rows, zero moments)     :905-977
//
// The reference does each of these as a chain of boolean-mask gathers, torch.cat and optimizer-state
// surgery: ~8 launches per tensor per operation over 7 parameter groups x 3 tensors + 6 statistics.
// Here every operation is ONE launch over a table of tensors passed by value:
//   adam_step_kernel        all groups, one launch (exp_avg / exp_avg_sq / param updated in one pass)
//   densify_stats_kernel    the five accumulators + max_radii2D of one view in one pass
//   densify_select_kernel   clone / split / prune decisions from the statistics in one pass
//   mask_* kernels          mask -> ascending row index list (block counts, scan, ordered scatter)
//   rows_gather_kernel      row gather of up to 32 tensors (parameters, moments, statistics) in one launch
//   split_rows_kernel       the new centres / scales of densify_and_split
// All of it is HBM-bound elementwise / gather work: no tensor cores.
#include <math.h>
#include <mutex>
#include "common.cuh"

namespace soar {

// ---- fused multi-tensor Adam ---------------------------------------------------------------------
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_ITEMS = 4;  // floats per thread

struct AdamSeg {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  long long numel;
  float step_size;  // lr / (1 - beta1^t), formed in double on the host like torch's python scalars
  float bc2_sqrt;   // sqrt(1 - beta2^t)
};
struct AdamTable {
  AdamSeg g[SOAR_ADAM_MAX_GROUPS];
  int block_start[SOAR_ADAM_MAX_GROUPS + 1];  // first CTA of each group
  int n;
  float w1, w2, beta2, eps;  // float(1 - beta1), float(1 - beta2), float(beta2), float(eps)
};

// torch/optim/adam.py _single_tensor_adam (no amsgrad, no weight decay, not maximize):
//   exp_avg.lerp_(grad, 1 - beta1)                       -> m + (g - m) * (1 - beta1)
//   exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2)
//   denom = (exp_avg_sq.sqrt() / sqrt(1 - beta2^t)).add_(eps)
//   param.addcdiv_(exp_avg, denom, value=-(lr / (1 - beta1^t)))
__global__ void __launch_bounds__(ADAM_THREADS) adam_step_kernel(const AdamTable t) {
  int gi = 0;
#pragma unroll 1
  while (gi + 1 < t.n && (int)blockIdx.x >= t.block_start[gi + 1]) ++gi;
  const AdamSeg g = t.g[gi];
  const long long base = ((long long)(blockIdx.x - t.block_start[gi]) * ADAM_THREADS + threadIdx.x) * ADAM_ITEMS;
  if (base >= g.numel) return;
  const float w1 = t.w1, w2 = t.w2, step_size = g.step_size, bc2_sqrt = g.bc2_sqrt;
  const bool vec = base + ADAM_ITEMS <= g.numel &&
                   ((((uintptr_t)g.param | (uintptr_t)g.grad | (uintptr_t)g.exp_avg | (uintptr_t)g.exp_avg_sq) & 15u) == 0);
  float p[ADAM_ITEMS], gr[ADAM_ITEMS], m[ADAM_ITEMS], v[ADAM_ITEMS];
  const int cnt = (int)min((long long)ADAM_ITEMS, g.numel - base);
  if (vec) {
    *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(g.param + base);
    *reinterpret_cast<float4*>(gr) = __ldg(reinterpret_cast<const float4*>(g.grad + base));
    *reinterpret_cast<float4*>(m) = *reinterpret_cast<const float4*>(g.exp_avg + base);
    *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(g.exp_avg_sq + base);
  } else {
    for (int i = 0; i < cnt; ++i) {
      p[i] = g.param[base + i]; gr[i] = g.grad[base + i]; m[i] = g.exp_avg[base + i]; v[i] = g.exp_avg_sq[base + i];
    }
  }
#pragma unroll
  for (int i = 0; i < ADAM_ITEMS; ++i) {
    if (i < cnt) {
      m[i] = m[i] + (gr[i] - m[i]) * w1;
      v[i] = v[i] * t.beta2 + (gr[i] * gr[i]) * w2;
      const float denom = sqrtf(v[i]) / bc2_sqrt + t.eps;
      p[i] = p[i] - step_size * (m[i] / denom);
    }
  }
  if (vec) {
    *reinterpret_cast<float4*>(g.param + base) = *reinterpret_cast<float4*>(p);
    *reinterpret_cast<float4*>(g.exp_avg + base) = *reinterpret_cast<float4*>(m);
    *reinterpret_cast<float4*>(g.exp_avg_sq + base) = *reinterpret_cast<float4*>(v);
  } else {
    for (int i = 0; i < cnt; ++i) {
      g.param[base + i] = p[i]; g.exp_avg[base + i] = m[i]; g.exp_avg_sq[base + i] = v[i];
    }
  }
}

// ---- per-view densification statistics ------------------------------------------------------------
// add_densification_stats(viewspace_point_tensor, update_filter) with update_filter = radii > 0, plus
// max_radii2D = max(max_radii2D, radii.float()) over ALL Gaussians (update_states :1214).
__global__ void __launch_bounds__(256)
densify_stats_kernel(int N, const float* __restrict__ viewspace_grad /*[N,3]*/,
                     const float* __restrict__ scaling_grad /*[N,3]*/,
                     const float* __restrict__ rotation /*[N,4]*/, const float* __restrict__ opacity /*[N]*/,
                     const int* __restrict__ radii, SoarDensifyStats s) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int r = radii[i];
  s.max_radii2D[i] = fmaxf(s.max_radii2D[i], (float)r);
  if (r <= 0) return;
  const float gx = viewspace_grad[3 * i], gy = viewspace_grad[3 * i + 1];
  s.xyz_gradient_accum[i] += sqrtf(gx * gx + gy * gy);
  s.scale_gradient_accum[i] += scaling_grad[3 * i] + scaling_grad[3 * i + 1];
  const float4 q = reinterpret_cast<const float4*>(rotation)[i];
  s.rot_gradient_accum[i] += sqrtf(((q.x * q.x + q.y * q.y) + q.z * q.z) + q.w * q.w);
  s.opac_gradient_accum[i] += opacity[i];
  s.denom[i] += 1.f;
}

// ---- selection rules ---------------------------------------------------------------------------------
__device__ __forceinline__ float nan_to_zero(float x) { return x != x ? 0.f : x; }

// mode 0: adaptive_densify -> out_a = clone mask, out_b = split mask
// mode 1: adaptive_prune   -> out_a = prune mask
__global__ void __launch_bounds__(256)
densify_select_kernel(int N, int mode, SoarDensifyStats s, const float* __restrict__ scaling /*[N,3] log*/,
                      const float* __restrict__ opacity /*[N] logit*/, float max_grad, float dense_extent,
                      float min_opacity, float half_extent, float area_thr,
                      uint8_t* __restrict__ out_a, uint8_t* __restrict__ out_b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float s0 = expf(scaling[3 * i]), s1 = expf(scaling[3 * i + 1]), s2 = expf(scaling[3 * i + 2]);
  const float den = s.denom[i];
  if (mode == 0) {
    const float grad_pos = nan_to_zero(s.xyz_gradient_accum[i] / den);
    const float grad_scale = nan_to_zero(s.scale_gradient_accum[i] / den);
    const float grad_opac = nan_to_zero(s.opac_gradient_accum[i] / den);
    const bool pre = (grad_opac <= 2.f) && (grad_scale <= 1e-7f);       // denser * larger (:1101-1103)
    const float smax = fmaxf(fmaxf(s0, s1), s2);
    // torch.norm of a [N,1] tensor over its last dim = |x|
    out_a[i] = (fabsf(grad_pos) >= max_grad) && (smax <= dense_extent) && pre;   // :1038-1046
    out_b[i] = (grad_pos >= max_grad) && (smax > dense_extent);                 // :987-994
  } else {
    const float op = 1.f / (1.f + expf(-opacity[i]));
    const float smin2 = fminf(s0, s1), smax2 = fmaxf(s0, s1);
    const bool prune_scale = (smax2 > half_extent) || ((smin2 * smax2) < area_thr);
    out_a[i] = (op < min_opacity) || (den == 0.f) || prune_scale;              // :1069-1081
  }
}

// ---- mask -> ascending index list -------------------------------------------------------------------
constexpr int MC_THREADS = 256;
constexpr int MC_ITEMS = 16;
constexpr int MC_TILE = MC_THREADS * MC_ITEMS;  // rows per CTA

__global__ void __launch_bounds__(MC_THREADS)
mask_count_kernel(const uint8_t* __restrict__ mask, int N, int invert, uint32_t* __restrict__ block_counts) {
  __shared__ uint32_t s_w[MC_THREADS / 32];
  const int base = blockIdx.x * MC_TILE;
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < MC_ITEMS; ++j) {
    const int i = base + j * MC_THREADS + threadIdx.x;
    if (i < N) c += ((mask[i] != 0) != (invert != 0)) ? 1u : 0u;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < MC_THREADS / 32; ++w) t += s_w[w];
    block_counts[blockIdx.x] = t;
  }
}

// exclusive scan of the block counts in place (one CTA), total -> *count
__global__ void __launch_bounds__(1024)
mask_scan_kernel(uint32_t* __restrict__ block_counts, int n_blocks, uint32_t* __restrict__ count) {
  __shared__ uint32_t s_scan[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + tid;
    const uint32_t c = i < n_blocks ? block_counts[i] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    uint32_t off = s_carry;
    for (int w = 0; w < warp; ++w) off += s_scan[w];
    if (i < n_blocks) block_counts[i] = off + incl - c;
    __syncthreads();
    if (tid == 1023) s_carry = off + incl;
    __syncthreads();
  }
  if (tid == 0) *count = s_carry;
}

__global__ void __launch_bounds__(MC_THREADS)
mask_scatter_kernel(const uint8_t* __restrict__ mask, int N, int invert,
                    const uint32_t* __restrict__ block_offsets, int32_t* __restrict__ index_out) {
  __shared__ uint32_t s_w[MC_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t running = block_offsets[blockIdx.x];
  const int base = blockIdx.x * MC_TILE;
#pragma unroll 1
  for (int j = 0; j < MC_ITEMS; ++j) {
    const int i = base + j * MC_THREADS + threadIdx.x;
    const bool sel = i < N && ((mask[i] != 0) != (invert != 0));
    const uint32_t bal = __ballot_sync(0xffffffffu, sel);
    if (lane == 0) s_w[warp] = __popc(bal);
    __syncthreads();
    uint32_t off = running, total = 0;
#pragma unroll
    for (int w = 0; w < MC_THREADS / 32; ++w) {
      const uint32_t c = s_w[w];
      if (w < warp) off += c;
      total += c;
    }
    if (sel) index_out[off + __popc(bal & ((1u << lane) - 1u))] = i;
    running += total;
    __syncthreads();
  }
}

// ---- multi-tensor row gather --------------------------------------------------------------------------
struct RowsTable {
  SoarRowTensor src[SOAR_ROWS_MAX_TENSORS];
  SoarRowTensor dst[SOAR_ROWS_MAX_TENSORS];
  int block_start[SOAR_ROWS_MAX_TENSORS + 1];
  int n;
};

// dst[t][dst_row0 + i, :] = src[t][index[i % n_index], :]  for i < n_out   (src NULL -> zeros)
__global__ void __launch_bounds__(256)
rows_gather_kernel(const RowsTable t, const int32_t* __restrict__ index, int n_index, long long n_out,
                   long long dst_row0) {
  int ti = 0;
#pragma unroll 1
  while (ti + 1 < t.n && (int)blockIdx.x >= t.block_start[ti + 1]) ++ti;
  const SoarRowTensor src = t.src[ti], dst = t.dst[ti];
  const int w = dst.width;
  const long long e = (long long)(blockIdx.x - t.block_start[ti]) * 256 + threadIdx.x;
  if (e >= n_out * w) return;
  const long long row = e / w;
  const int col = (int)(e - row * w);
  float v = 0.f;
  if (src.data != nullptr) {
    const long long srow = index ? (long long)index[row % n_index] : row;
    v = src.data[srow * w + col];
  }
  dst.data[(dst_row0 + row) * w + col] = v;
}

// ---- densify_and_split: the new centres and scales (:996-1010) ------------------------------------------
// For i < n_sel * n_rep, j = index[i % n_sel]:
//   xyz' = R(normalize(rotation[j])) . samples[i] + xyz[j]
//   scaling' = log(exp(scaling[j]) / (0.8 * n_rep)); scaling'.z = -1e10 for surfels (config[0] > 0)
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ xyz, const float* __restrict__ scaling,
                  const float* __restrict__ rotation, const int32_t* __restrict__ index, int n_sel,
                  int n_rep, const float* __restrict__ samples, int surfel, float* __restrict__ out_xyz,
                  float* __restrict__ out_scaling) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sel * n_rep) return;
  const int j = index[i % n_sel];
  const float4 q4 = reinterpret_cast<const float4*>(rotation)[j];
  const float norm = sqrtf(q4.x * q4.x + q4.y * q4.y + q4.z * q4.z + q4.w * q4.w);
  const float r = q4.x / norm, x = q4.y / norm, y = q4.z / norm, z = q4.w / norm;
  const float sx = samples[3 * i], sy = samples[3 * i + 1], sz = samples[3 * i + 2];
  const float R00 = 1 - 2 * (y * y + z * z), R01 = 2 * (x * y - r * z), R02 = 2 * (x * z + r * y);
  const float R10 = 2 * (x * y + r * z), R11 = 1 - 2 * (x * x + z * z), R12 = 2 * (y * z - r * x);
  const float R20 = 2 * (x * z - r * y), R21 = 2 * (y * z + r * x), R22 = 1 - 2 * (x * x + y * y);
  out_xyz[3 * i + 0] = (R00 * sx + R01 * sy + R02 * sz) + xyz[3 * j + 0];
  out_xyz[3 * i + 1] = (R10 * sx + R11 * sy + R12 * sz) + xyz[3 * j + 1];
  out_xyz[3 * i + 2] = (R20 * sx + R21 * sy + R22 * sz) + xyz[3 * j + 2];
  const float div = 0.8f * (float)n_rep;
#pragma unroll
  for (int a = 0; a < 3; ++a) out_scaling[3 * i + a] = logf(expf(scaling[3 * j + a]) / div);
  if (surfel) out_scaling[3 * i + 2] = -1e10f;
}

}  // namespace soar

using namespace soar;

extern "C" {

int soar_adam_step(const SoarAdamGroup* groups, int32_t n_groups, double beta1, double beta2,
                   double eps, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_groups < 0 || n_groups > SOAR_ADAM_MAX_GROUPS || (n_groups > 0 && !groups)) {
    set_error("soar_adam_step: %d groups (0..%d)", n_groups, SOAR_ADAM_MAX_GROUPS);
    return SOAR_E_INVALID;
  }
  AdamTable t;
  t.n = 0;
  t.w1 = (float)(1.0 - beta1); t.w2 = (float)(1.0 - beta2); t.beta2 = (float)beta2; t.eps = (float)eps;
  int blocks = 0;
  for (int i = 0; i < n_groups; ++i) {
    const SoarAdamGroup& g = groups[i];
    if (g.numel < 0 || g.step < 1) {
      set_error("soar_adam_step: group %d has numel %lld, step %d", i, (long long)g.numel, g.step);
      return SOAR_E_INVALID;
    }
    if (g.numel == 0) continue;
    if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) {
      set_error("soar_adam_step: group %d has a NULL pointer", i);
      return SOAR_E_INVALID;
    }
    AdamSeg& a = t.g[t.n];
    a.param = g.param; a.grad = g.grad; a.exp_avg = g.exp_avg; a.exp_avg_sq = g.exp_avg_sq;
    a.numel = g.numel;
    a.step_size = (float)(g.lr / (1.0 - pow(beta1, (double)g.step)));
    a.bc2_sqrt = (float)sqrt(1.0 - pow(beta2, (double)g.step));
    t.block_start[t.n] = blocks;
    blocks += (int)((g.numel + ADAM_THREADS * ADAM_ITEMS - 1) / (ADAM_THREADS * ADAM_ITEMS));
    ++t.n;
  }
  t.block_start[t.n] = blocks;
  if (blocks == 0) return SOAR_OK;
  {
    KernelScope ks(K_ADAM_STEP, stream);
    adam_step_kernel<<<blocks, ADAM_THREADS, 0, stream>>>(t);
  }
  SOAR_LAUNCH_CHECK("adam_step_kernel");
  return SOAR_OK;
}

static bool stats_ok(const SoarDensifyStats* s) {
  return s && s->xyz_gradient_accum && s->scale_gradient_accum && s->rot_gradient_accum &&
         s->opac_gradient_accum && s->denom && s->max_radii2D;
}

int soar_densify_add_stats(int32_t N, const float* viewspace_grad, const float* scaling_grad,
                           const float* rotation, const float* opacity, const int32_t* radii,
                           const SoarDensifyStats* stats, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0) { set_error("soar_densify_add_stats: N < 0"); return SOAR_E_INVALID; }
  if (N == 0) return SOAR_OK;
  if (!viewspace_grad || !scaling_grad || !rotation || !opacity || !radii || !stats_ok(stats)) {
    set_error("soar_densify_add_stats: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  {
    KernelScope ks(K_DENSIFY_STATS, stream);
    densify_stats_kernel<<<ceil_div(N, 256), 256, 0, stream>>>(N, viewspace_grad, scaling_grad, rotation,
                                                               opacity, radii, *stats);
  }
  SOAR_LAUNCH_CHECK("densify_stats_kernel");
  return SOAR_OK;
}

int soar_densify_select(int32_t N, int32_t mode, const SoarDensifyStats* stats, const float* scaling,
                        const float* opacity, float max_grad, float dense_extent, float min_opacity,
                        float half_extent, float area_thr, uint8_t* out_a, uint8_t* out_b,
                        soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0 || (mode != 0 && mode != 1)) { set_error("soar_densify_select: bad N/mode"); return SOAR_E_INVALID; }
  if (N == 0) return SOAR_OK;
  if (!stats_ok(stats) || !scaling || !out_a || (mode == 0 && !out_b) || (mode == 1 && !opacity)) {
    set_error("soar_densify_select: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  {
    KernelScope ks(K_DENSIFY_SELECT, stream);
    densify_select_kernel<<<ceil_div(N, 256), 256, 0, stream>>>(
        N, mode, *stats, scaling, opacity, max_grad, dense_extent, min_opacity, half_extent, area_thr,
        out_a, out_b);
  }
  SOAR_LAUNCH_CHECK("densify_select_kernel");
  return SOAR_OK;
}

size_t soar_mask_index_workspace_bytes(int32_t N) {
  return (size_t)(ceil_div(N < 0 ? 0 : N, MC_TILE) + 1) * sizeof(uint32_t);
}

int soar_mask_to_index(const uint8_t* mask, int32_t N, int32_t invert, int32_t* index_out,
                       uint32_t* count_out, void* workspace, size_t workspace_bytes,
                       soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0 || !count_out) { set_error("soar_mask_to_index: bad arguments"); return SOAR_E_INVALID; }
  if (N == 0) {
    SOAR_CUDA_TRY(cudaMemsetAsync(count_out, 0, sizeof(uint32_t), stream));
    return SOAR_OK;
  }
  if (!mask || !index_out || !workspace || workspace_bytes < soar_mask_index_workspace_bytes(N)) {
    set_error("soar_mask_to_index: NULL pointer or workspace too small");
    return SOAR_E_WORKSPACE;
  }
  const int blocks = ceil_div(N, MC_TILE);
  uint32_t* counts = reinterpret_cast<uint32_t*>(workspace);
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_count_kernel<<<blocks, MC_THREADS, 0, stream>>>(mask, N, invert, counts);
  }
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_scan_kernel<<<1, 1024, 0, stream>>>(counts, blocks, count_out);
  }
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_scatter_kernel<<<blocks, MC_THREADS, 0, stream>>>(mask, N, invert, counts, index_out);
  }
  SOAR_LAUNCH_CHECK("mask_to_index kernels");
  return SOAR_OK;
}

// ---- soar_mask_to_index for callers that must not allocate -----------------------------------------
// The per-CTA counts live in this array, which is part of the library image: every GPU that loads
// the library carries SOAR_MASK_INDEX_STATIC_MAX_ROWS / MC_TILE + 1 words (16 KB) of it until the
// process exits. On the host the NAME `g_mask_static_counts` is only a handle for the runtime (its
// value is not a device address and must never be handed to a kernel); the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int MASK_STATIC_MAX_BLOCKS = SOAR_MASK_INDEX_STATIC_MAX_ROWS / MC_TILE;
static_assert(SOAR_MASK_INDEX_STATIC_MAX_ROWS % MC_TILE == 0, "the limit is a whole number of CTA tiles");
static __device__ __align__(256) uint32_t g_mask_static_counts[MASK_STATIC_MAX_BLOCKS + 1];

constexpr int MASK_STATIC_MAX_DEVICES = 64;
static uint32_t* g_mask_static_ptr[MASK_STATIC_MAX_DEVICES];  // per device ordinal, NULL = not asked yet
static std::mutex g_mask_static_mutex;

// Device address of this device's copy of the array (set-up on the first call per device).
static int mask_static_counts(uint32_t** out) {
  int dev = -1;
  SOAR_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MASK_STATIC_MAX_DEVICES) {
    set_error("soar_mask_to_index_static: device ordinal %d outside 0..%d", dev, MASK_STATIC_MAX_DEVICES - 1);
    return SOAR_E_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_mask_static_mutex);
  if (!g_mask_static_ptr[dev]) {
    void* p = nullptr;
    SOAR_CUDA_TRY(cudaGetSymbolAddress(&p, g_mask_static_counts));
    if (!p) {
      set_error("soar_mask_to_index_static: cudaGetSymbolAddress gave a NULL address on device %d", dev);
      return SOAR_E_CUDA;
    }
    g_mask_static_ptr[dev] = static_cast<uint32_t*>(p);
  }
  *out = g_mask_static_ptr[dev];
  return SOAR_OK;
}

// Same three launches as soar_mask_to_index; only the home of the per-CTA counts differs. Nothing in
// the array has to start at zero: the count kernel stores (not adds) words 0..blocks-1, the scan
// rewrites exactly those, the scatter reads exactly those — so whatever an earlier call (a longer
// mask, say) left behind is never read.
int soar_mask_to_index_static(const uint8_t* mask, int32_t N, int32_t invert, int32_t* index_out,
                              uint32_t* count_out, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0 || !count_out) { set_error("soar_mask_to_index_static: bad arguments"); return SOAR_E_INVALID; }
  if (N > SOAR_MASK_INDEX_STATIC_MAX_ROWS) {
    set_error("soar_mask_to_index_static: %d rows, the built-in scratch holds at most %d "
              "(use soar_mask_to_index with a caller-owned workspace)", N, SOAR_MASK_INDEX_STATIC_MAX_ROWS);
    return SOAR_E_WORKSPACE;
  }
  if (N == 0) {
    SOAR_CUDA_TRY(cudaMemsetAsync(count_out, 0, sizeof(uint32_t), stream));
    return SOAR_OK;
  }
  if (!mask || !index_out) {
    set_error("soar_mask_to_index_static: NULL mask or index_out");
    return SOAR_E_INVALID;
  }
  uint32_t* counts = nullptr;
  const int rc = mask_static_counts(&counts);
  if (rc != SOAR_OK) return rc;
  const int blocks = ceil_div(N, MC_TILE);
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_count_kernel<<<blocks, MC_THREADS, 0, stream>>>(mask, N, invert, counts);
  }
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_scan_kernel<<<1, 1024, 0, stream>>>(counts, blocks, count_out);
  }
  {
    KernelScope ks(K_MASK_INDEX, stream);
    mask_scatter_kernel<<<blocks, MC_THREADS, 0, stream>>>(mask, N, invert, counts, index_out);
  }
  SOAR_LAUNCH_CHECK("mask_to_index_static kernels");
  return SOAR_OK;
}

int soar_rows_gather(const SoarRowTensor* src, const SoarRowTensor* dst, int32_t n_tensors,
                     const int32_t* index, int32_t n_index, int64_t n_out, int64_t dst_row0,
                     soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_tensors < 0 || n_tensors > SOAR_ROWS_MAX_TENSORS || n_out < 0 || dst_row0 < 0 ||
      (n_tensors > 0 && (!src || !dst)) || (index && n_index <= 0 && n_out > 0)) {
    set_error("soar_rows_gather: bad arguments (%d tensors, max %d)", n_tensors, SOAR_ROWS_MAX_TENSORS);
    return SOAR_E_INVALID;
  }
  if (n_out == 0 || n_tensors == 0) return SOAR_OK;
  RowsTable t;
  t.n = 0;
  long long blocks = 0;
  for (int i = 0; i < n_tensors; ++i) {
    if (dst[i].width <= 0 || src[i].width != dst[i].width || !dst[i].data) {
      set_error("soar_rows_gather: tensor %d has width %d/%d or a NULL destination", i, src[i].width, dst[i].width);
      return SOAR_E_INVALID;
    }
    t.src[t.n] = src[i]; t.dst[t.n] = dst[i];
    t.block_start[t.n] = (int)blocks;
    blocks += (n_out * dst[i].width + 255) / 256;
    ++t.n;
  }
  if (blocks > 0x7fffffffLL) { set_error("soar_rows_gather: too many elements"); return SOAR_E_INVALID; }
  t.block_start[t.n] = (int)blocks;
  {
    KernelScope ks(K_ROWS_GATHER, stream);
    rows_gather_kernel<<<(int)blocks, 256, 0, stream>>>(t, index, n_index, n_out, dst_row0);
  }
  SOAR_LAUNCH_CHECK("rows_gather_kernel");
  return SOAR_OK;
}

int soar_densify_split_rows(const float* xyz, const float* scaling, const float* rotation,
                            const int32_t* index, int32_t n_sel, int32_t n_rep, const float* samples,
                            int32_t surfel, float* out_xyz, float* out_scaling, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_sel < 0 || n_rep <= 0) { set_error("soar_densify_split_rows: bad counts"); return SOAR_E_INVALID; }
  if (n_sel == 0) return SOAR_OK;
  if (!xyz || !scaling || !rotation || !index || !samples || !out_xyz || !out_scaling) {
    set_error("soar_densify_split_rows: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  {
    KernelScope ks(K_SPLIT_ROWS, stream);
    split_rows_kernel<<<ceil_div(n_sel * n_rep, 256), 256, 0, stream>>>(
        xyz, scaling, rotation, index, n_sel, n_rep, samples, surfel, out_xyz, out_scaling);
  }
  SOAR_LAUNCH_CHECK("split_rows_kernel");
  return SOAR_OK;
}

}  // extern "C"
