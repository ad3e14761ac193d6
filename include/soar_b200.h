/*
 * soar_b200.h — C ABI of the B200-native SOAR render-loop kernels.
 *
 * This is the drop-in boundary. Every entry point names the
 * reference interface it replaces; paths are relative to the reference tree
 * (DGR = submodules/diff-gaussian-rasterization, KNN = submodules/simple-knn,
 * TS = soar/threestudio-soar).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types. All pointers are DEVICE
 *     pointers to contiguous float32/int32 unless the name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - the library keeps no global state besides a thread-local error string;
 *     every scratch buffer is caller-owned (the python shim allocates them with
 *     torch and keeps them in the autograd ctx, like the reference does through
 *     its resize_ callbacks, DGR/rasterize_points.cu:27-33). The single
 *     exception, soar_mask_to_index_static, says so at its declaration.
 *   - return 0 on success, negative SOAR_E_* on error; soar_last_error() gives
 *     the message (the shim re-raises it as RuntimeError like the reference's
 *     C++ exceptions, DGR/rasterize_points.cu:50-52).
 */
#ifndef SOAR_B200_H_INCLUDED
#define SOAR_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SOAR_OK 0
#define SOAR_E_INVALID (-1)  /* bad argument (shape, null pointer, size)        */
#define SOAR_E_CUDA (-2)     /* a CUDA runtime call or launch failed            */
#define SOAR_E_WORKSPACE (-3) /* caller-provided scratch buffer too small       */
#define SOAR_E_PREFILTER (-4) /* prefiltered=1 but a Gaussian was culled
                                 (reference: device __trap, DGR auxiliary.h:163) */

typedef void* soar_stream_t;

/* One view = GaussianRasterizationSettings + the per-call tensors
 * (DGR/diff_gaussian_rasterization/__init__.py:267-284 and the 24-argument
 * tuple at :77-102 <-> DGR/rasterize_points.h:18-32). Matrices are in the
 * reference's row-vector convention, i.e. read column-major by the kernels
 * (DGR/cuda_rasterizer/auxiliary.h:65-84). */
typedef struct SoarRastView {
  int32_t P;          /* number of Gaussians                                    */
  int32_t sh_degree;  /* active SH degree D (0..3)                              */
  int32_t M;          /* SH coefficients per Gaussian (0 when colors_precomp)   */
  int32_t width, height;
  float tanfovx, tanfovy, scale_modifier;
  int32_t prefiltered, render_front, sort_descending, debug;
  const float* bg;             /* [3]                                            */
  const float* means3D;        /* [P,3]                                          */
  const float* shs;            /* [P,M,3] or NULL                                */
  const float* colors_precomp; /* [P,3]   or NULL                                */
  const float* opacities;      /* [P]                                            */
  const float* scales;         /* [P,3]   or NULL                                */
  const float* rotations;      /* [P,4] (w,x,y,z), NOT normalised by the kernel;
                                  16-byte aligned -> vector loads, else scalar   */
  const float* cov3D_precomp;  /* [P,6]   or NULL                                */
  const float* viewmatrix;     /* [16]                                           */
  const float* projmatrix;     /* [16]                                           */
  const float* prcppoint;      /* [2]                                            */
  const float* patch_bbox;     /* [4] = (h0,w0,h1,w1)                            */
  const float* campos;         /* [3]                                            */
  const float* config;         /* [4] = (surface,normalize_depth,perpix_depth,lrn_cam) */
  /* Extension: what the HOST knows about `config` (the reference only has the device tensor,
   * DGR/cuda_rasterizer/forward.cu:464, backward.cu:581). SOAR_CFG_KNOWN set = the other bits equal
   * `config[i] > 0` and the launchers pick the specialised blend kernels on the host; 0 = unknown, the
   * kernels branch on the device copy. A set bit that contradicts the device tensor is the caller's
   * error (the device copy is still what preprocess and the geometry backward read). */
  uint32_t config_host_flags;
} SoarRastView;
#define SOAR_CFG_KNOWN           (1u << 0)
#define SOAR_CFG_SURFACE         (1u << 1)
#define SOAR_CFG_NORMALIZE_DEPTH (1u << 2)
#define SOAR_CFG_PERPIX_DEPTH    (1u << 3)
#define SOAR_CFG_LRN_CAM         (1u << 4)

/* Gradient outputs of the rasterizer backward: the 11-tuple returned by
 * RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:107-187). The kernels
 * write EVERY element (zeros for culled Gaussians), so the caller may pass
 * uninitialised memory — the reference needs 15 torch::zeros fills (:133-147). */
typedef struct SoarRastGrads {
  float* dL_dmeans2D;   /* [P,3]  NDC-scaled screen gradient, .z = 0            */
  float* dL_dcolors;    /* [P,3]                                                */
  float* dL_dopacity;   /* [P,1]                                                */
  float* dL_dmeans3D;   /* [P,3]                                                */
  float* dL_dcov3D;     /* [P,6]                                                */
  float* dL_dsh;        /* [P,M,3] (may be NULL when M == 0)                    */
  float* dL_dscales;    /* [P,3]                                                */
  float* dL_drotations; /* [P,4]  (16-byte aligned -> vector stores, else scalar)  */
  float* dL_dviewmat;   /* [16]  nonzero only when config[3] > 0                */
  float* dL_dprojmat;   /* [16]                                                 */
  float* dL_dcampos;    /* [3]                                                  */
  /* Extension: bit SOAR_ACC_* set = ADD this view's gradient to what the buffer already holds
   * (culled Gaussians then write nothing) instead of overwriting it. 0 = reference behaviour. */
  uint32_t accumulate_mask;
} SoarRastGrads;
#define SOAR_ACC_MEANS2D   (1u << 0)
#define SOAR_ACC_COLORS    (1u << 1)
#define SOAR_ACC_OPACITY   (1u << 2)
#define SOAR_ACC_MEANS3D   (1u << 3)
#define SOAR_ACC_COV3D     (1u << 4)
#define SOAR_ACC_SH        (1u << 5)
#define SOAR_ACC_SCALES    (1u << 6)
#define SOAR_ACC_ROTATIONS (1u << 7)

/* ---- rasterizer (replaces DGR `_C.rasterize_gaussians`, DGR/ext.cpp:16 ->
 *      RasterizeGaussiansCUDA -> CudaRasterizer::Rasterizer::forward,
 *      DGR/cuda_rasterizer/rasterizer_impl.cu:188-312) -------------------------- */

/* Sizes of the caller-owned scratch buffers (reference: required<State>(),
 * DGR/cuda_rasterizer/rasterizer_impl.h:74-84). */
size_t soar_rast_geom_bytes(int32_t P);
size_t soar_rast_image_bytes(int32_t width, int32_t height);
size_t soar_rast_binning_bytes(uint32_t num_rendered, int32_t width, int32_t height);

/* Phase 1 (K1+K2): per-Gaussian preprocess, depth order of the Gaussians
 * (onesweep radix sort of P depth keys) + decoupled-look-back prefix sum of
 * tiles_touched in that order. Writes radii[P], fills `geom`, and returns num_rendered
 * through *num_rendered_host after synchronising `stream` (the reference blocks
 * at the same point, rasterizer_impl.cu:250). */
int soar_rast_forward_bin(const SoarRastView* view, void* geom, size_t geom_bytes,
                          int32_t* radii, uint32_t* num_rendered_host, soar_stream_t stream);

/* Same as soar_rast_forward_bin without the host synchronisation: the count (and, when
 * view->prefiltered, the error flag in word 1) are copied asynchronously into
 * status_pinned_host[2], which must be page-locked host memory that stays valid until `stream`
 * has passed this point. A multi-view step enqueues phase 1 of ALL its views first and only
 * then waits for the first count, so the device never idles behind the host. The caller checks
 * word 1 (non-zero = SOAR_E_PREFILTER) after synchronising. */
int soar_rast_forward_bin_async(const SoarRastView* view, void* geom, size_t geom_bytes,
                                int32_t* radii, uint32_t* status_pinned_host,
                                soar_stream_t stream);

/* Phase 2 (K3..K6): (tile, Gaussian) pair emission in depth order, stable
 * onesweep radix sort over the tile bits, tile ranges, tile alpha-blend. The
 * resulting list order equals the reference's sort of tile|depth keys. Outputs: color[3,H,W] normal[3,H,W]
 * depth[1,H,W] opac[1,H,W]; every pixel is written. */
int soar_rast_forward_render(const SoarRastView* view, void* geom, void* binning,
                             size_t binning_bytes, void* img, size_t img_bytes,
                             uint32_t num_rendered, float* out_color, float* out_normal,
                             float* out_depth, float* out_opac, soar_stream_t stream);

/* Backward (replaces `_C.rasterize_gaussians_backward`, DGR/ext.cpp:17 ->
 * Rasterizer::backward, rasterizer_impl.cu:316-379). `geom`, `binning`, `img`
 * are the buffers filled by the forward of the same view. `scratch` must hold
 * soar_rast_backward_scratch_bytes(P) bytes. */
size_t soar_rast_backward_scratch_bytes(int32_t P);
int soar_rast_backward(const SoarRastView* view, const int32_t* radii, const void* geom,
                       const void* binning, const void* img, uint32_t num_rendered,
                       const float* dL_dout_color, const float* dL_dout_normal,
                       const float* dL_dout_depth, const float* dL_dout_opac,
                       void* scratch, size_t scratch_bytes, const SoarRastGrads* grads,
                       soar_stream_t stream);

/* Extension for view loops that run several views on several streams and accumulate into the
 * same gradient buffers (SoarRastGrads.accumulate_mask): the same backward, but the launch that
 * writes the gradient tensors waits for `accumulate_after` (a cudaEvent_t, NULL = no wait) and
 * `accumulated` (a cudaEvent_t, NULL = none) is recorded behind it. Chaining the views of a step
 * through these events keeps the order of the additions — and so the rounding — fixed, and lets
 * the first view of a step store instead of add, while the blend replays of different views
 * still overlap. */
typedef void* soar_event_t;
int soar_rast_backward_ordered(const SoarRastView* view, const int32_t* radii, const void* geom,
                               const void* binning, const void* img, uint32_t num_rendered,
                               const float* dL_dout_color, const float* dL_dout_normal,
                               const float* dL_dout_depth, const float* dL_dout_opac,
                               void* scratch, size_t scratch_bytes, const SoarRastGrads* grads,
                               soar_event_t accumulate_after, soar_event_t accumulated,
                               soar_stream_t stream);

/* Replaces `_C.mark_visible` (DGR/ext.cpp:18). The reference kernel body is
 * commented out (rasterizer_impl.cu:52-62): `present` is left untouched, so the
 * all-false tensor the caller allocated is what comes back. */
int soar_rast_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                           const float* projmatrix, uint8_t* present, soar_stream_t stream);

/* Debug accessors for parity tests: device pointers into the
 * private scratch layout. what: 0 sorted tile ids u32[R], 1 point_list u32[R],
 * 2 ranges uint2[tiles], 3 n_contrib u32[W*H], 4 final_T f32[W*H],
 * 5 tiles_touched u32[P], 6 point_offsets u32[P] (inclusive, in depth order),
 * 7 records f32[P*24], 8 final_D f32[W*H], 9 depth-ordered Gaussian indices u32[P]. */
const void* soar_rast_debug_ptr(int what, int32_t P, int32_t width, int32_t height,
                                uint32_t num_rendered, const void* geom, const void* binning,
                                const void* img);
/* Materialises the reference's sorted 64-bit keys (tile << 32 | depth bits,
 * DGR/cuda_rasterizer/rasterizer_impl.cu:66-99 after the sort at :266-285) into
 * keys_out[R] (device). The product never builds them: it sorts Gaussians by
 * depth once and the (tile, index) pairs by tile only. */
int soar_rast_debug_sorted_keys(int32_t P, int32_t width, int32_t height, uint32_t num_rendered,
                                const void* geom, const void* binning, uint64_t* keys_out,
                                soar_stream_t stream);

/* ---- KNN (replaces `simple_knn._C.distCUDA2`, KNN/ext.cpp:15-17 ->
 *      SimpleKNN::knn, KNN/simple_knn.cu:186-222) ------------------------------- */
size_t soar_knn_workspace_bytes(int32_t P);
int soar_knn_dist2(const float* points, int32_t P, float* mean_dist2, void* workspace,
                   size_t workspace_bytes, soar_stream_t stream);

/* ---- LBS skinning (replaces the torch op chain at
 *      TS/utils/smpl.py:613 (weights x bones blend) and
 *      TS/renderer/diff_gaussian_rasterizer.py:103-114 (apply to points/quats)) -- */
/* weights[N,J] (row-major), bones[J,16] row-major 4x4 (cano->live), points[N,3],
 * quats[N,4] (w,x,y,z). quats, out_quats, out_mats, mats and the quaternion gradients are moved as
 * 16-byte rows: their base pointers must be 16-byte aligned (SOAR_E_INVALID otherwise).
 * axis_perm: 0 = none (gt views, :138-149),
 * 1 = the SDS-view swizzle (x,y,z)->(z,x,y) of transform_point_cloud("+z,+x,+y"). */
int soar_lbs_forward(const float* weights, const float* bones, int32_t J, const float* points,
                     const float* quats, int32_t N, int32_t axis_perm, float* out_points,
                     float* out_quats, float* out_mats /* [N,12] or NULL */,
                     soar_stream_t stream);
/* `mats` is the [N,12] per-point transform the forward wrote to out_mats (the
 * backward reads 48 B/point instead of re-blending the 220-byte weight row). */
int soar_lbs_backward(const float* mats, const float* quats, int32_t N, int32_t axis_perm,
                      const float* dL_dout_points, const float* dL_dout_quats,
                      float* dL_dpoints, float* dL_dquats, soar_stream_t stream);

/* Skin-weight query (replaces SMPL_Guidance.query_weights_smpl, TS/utils/smpl.py:618-637:
 * pytorch3d.knn_points(K=30) + sqrt/clamp(1e-4, 1) + inverse-distance blend of the K nearest
 * canonical vertices' skin-weight rows), optionally fused with the bone blend of :613.
 * points[N,3], cano_vertices[V,3] (V <= 16384), lbs_weights[V,J] (J <= 64), K <= 32.
 * out_weights[N,J] and/or out_mats[N,12] (needs bones[J,16]); either may be NULL, not both. */
int soar_lbs_query_weights(const float* points, int32_t N, const float* cano_vertices, int32_t V,
                           const float* lbs_weights, int32_t J, int32_t K, const float* bones,
                           float* out_weights, float* out_mats, soar_stream_t stream);

/* The same query through a uniform grid over the canonical vertices (exact: identical neighbours,
 * ties on the vertex index): a few hundred distance evaluations per point instead of V. The grid
 * is built once per vertex set (they change only when betas change, TS/utils/smpl.py:249-258). */
size_t soar_lbs_query_grid_bytes(int32_t V);
int soar_lbs_query_build_grid(const float* cano_vertices, int32_t V, void* grid, size_t grid_bytes,
                              soar_stream_t stream);
int soar_lbs_query_weights_grid(const float* points, int32_t N, const void* grid, int32_t V,
                                const float* lbs_weights, int32_t J, int32_t K, const float* bones,
                                float* out_weights, float* out_mats, soar_stream_t stream);

/* The same query, batched for large N (exact, bit-identical results): the queries are sorted by grid
 * cell with the library's own radix sort and a warp walks the search cube of a cell once for all its
 * queries (lbs_query.cu). `workspace` holds the sort's ping-pong buffers,
 * soar_lbs_query_cells_workspace_bytes(N) bytes. K != 30, V > 13000, small N or a NULL workspace
 * take soar_lbs_query_weights_grid's per-point walk. */
size_t soar_lbs_query_cells_workspace_bytes(int32_t N);
int soar_lbs_query_weights_cells(const float* points, int32_t N, const void* grid, int32_t V,
                                 const float* lbs_weights, int32_t J, int32_t K, const float* bones,
                                 float* out_weights, float* out_mats, void* workspace,
                                 size_t workspace_bytes, soar_stream_t stream);

/* ---- SMPL-X body-model forward for one pose (replaces smplx.lbs.lbs(),
 *      TS/utils/smplx/lbs.py:147-246, plus the transl fold of SMPLX.forward,
 *      TS/utils/smplx/body_models.py:1380-1383, and the cano->live bones of
 *      TS/utils/smpl.py:609) ---------------------------------------------------------- */
typedef struct SoarSmplxModel {
  int32_t V, J, NB;           /* vertices, joints (<= 64), shape coefficients (<= 64) */
  const float* v_template;    /* [V,3]                                                */
  const float* shapedirs;     /* [V,3,NB]                                             */
  const float* posedirs;      /* [(J-1)*9, V*3]                                       */
  const float* J_regressor;   /* [J,V]                                                */
  const int32_t* parents;     /* [J], parents[0] ignored                              */
  const float* lbs_weights;   /* [V,J]                                                */
} SoarSmplxModel;
size_t soar_smplx_workspace_bytes(int32_t V, int32_t J);
/* betas[NB], full_pose[J*3] axis-angle, transl[3] or NULL, inv_mats[J,16] or NULL.
 * Outputs (each may be NULL): out_verts[V,3], out_joints[J,3], out_A[J,16] with transl folded
 * in (Q17), out_bones[J,16] = out_A . inv_mats. */
int soar_smplx_forward(const SoarSmplxModel* model, const float* betas, const float* full_pose,
                       const float* transl, const float* inv_mats, float* out_verts,
                       float* out_joints, float* out_A, float* out_bones, void* workspace,
                       size_t workspace_bytes, soar_stream_t stream);

/* ---- image-space post-ops of the surfel renderer (replaces the torch op chains
 *      depth2normal / normal2curv, TS/renderer/diff_gaussian_rasterizer.py:359-448, and the
 *      renderer's post block at :292-303) -------------------------------------------------- */
typedef struct SoarPostArgs {
  int32_t width, height;
  float cx, cy;            /* prcppoint[0]*width, prcppoint[1]*height (:375-376)              */
  float kinv_x, kinv_y;    /* 1/fov2focal(FoVy, height), 1/fov2focal(FoVx, width) — the
                            * reference's K00/K11 pairing, kept (:378-382)                    */
  float mask_threshold;    /* mask = opac > threshold when `mask` is NULL (1e-5 at :292)      */
  int32_t surfel;          /* 0: the two functions as defined; 1: the renderer's post block:
                            * channels 1,2 negated, (x+1)/2, normal gradient cut off-mask     */
  const uint8_t* mask;     /* [H,W] bool, or NULL                                             */
  const float* opac;       /* [H,W], used when mask is NULL                                   */
  const float* normal;     /* [3,H,W] or NULL (no curvature / normal output)                  */
  const float* depth;      /* [H,W] or NULL (no depth normal)                                 */
} SoarPostArgs;
/* One launch. Outputs (each may be NULL): out_normal[3,H,W] (surfel only), out_curv[H,W],
 * out_depth_normal[3,H,W], out_mask[H,W] (bool). */
int soar_post_forward(const SoarPostArgs* args, float* out_normal, float* out_curv,
                      float* out_depth_normal, uint8_t* out_mask, soar_stream_t stream);
/* One launch, deterministic gather (no atomics). Incoming gradients may be NULL (= zero);
 * dL_dnormal[3,H,W] / dL_ddepth[H,W] are fully written when given. */
int soar_post_backward(const SoarPostArgs* args, const float* dL_dout_normal, const float* dL_dcurv,
                       const float* dL_ddepth_normal, float* dL_dnormal, float* dL_ddepth,
                       soar_stream_t stream);

/* ---- optimizer step and densify / clone / split / prune of the Gaussian parameter groups
 *      (replaces torch.optim.Adam over the per-Gaussian groups, TS/geometry/surfel_base.py:596-681,
 *      and the boolean-mask / torch.cat / optimizer-state surgery of :847-1135) ------------------ */
#define SOAR_ADAM_MAX_GROUPS 16
typedef struct SoarAdamGroup {
  float* param;              /* updated in place                                   */
  const float* grad;
  float* exp_avg;            /* Adam first moment, in place                        */
  float* exp_avg_sq;         /* Adam second moment, in place                       */
  int64_t numel;
  double lr;                 /* this group's learning rate for this step           */
  int32_t step;              /* 1-based step count of this group's state           */
  int32_t reserved;
} SoarAdamGroup;
/* One launch for all groups; torch.optim.Adam semantics (no amsgrad / weight decay). The
 * hyper-parameters are doubles like torch's python scalars: 1 - beta, the bias corrections and the
 * step size are formed in double and rounded to float once. */
int soar_adam_step(const SoarAdamGroup* groups, int32_t n_groups, double beta1, double beta2,
                   double eps, soar_stream_t stream);

typedef struct SoarDensifyStats {      /* each [N] (the reference keeps [N,1])      */
  float* xyz_gradient_accum;
  float* scale_gradient_accum;
  float* rot_gradient_accum;
  float* opac_gradient_accum;
  float* denom;
  float* max_radii2D;
} SoarDensifyStats;
/* add_densification_stats (:1113-1135) with update_filter = radii > 0, plus
 * max_radii2D = max(max_radii2D, radii) (:1214), for one rendered view. viewspace_grad [N,3],
 * scaling_grad [N,3], rotation [N,4], opacity [N] (the logit parameter), radii [N]. */
int soar_densify_add_stats(int32_t N, const float* viewspace_grad, const float* scaling_grad,
                           const float* rotation, const float* opacity, const int32_t* radii,
                           const SoarDensifyStats* stats, soar_stream_t stream);
/* mode 0 = adaptive_densify (:1083-1111): out_a = clone mask (:1038-1046), out_b = split mask
 * (:987-994); mode 1 = adaptive_prune (:1069-1081): out_a = prune mask. scaling [N,3] (log),
 * opacity [N] (logit). dense_extent = percent_dense * scene_extent, half_extent = 0.5 * extent,
 * area_thr = 1e-8 * extent^2 (formed by the caller in double, like the python scalars). */
int soar_densify_select(int32_t N, int32_t mode, const SoarDensifyStats* stats, const float* scaling,
                        const float* opacity, float max_grad, float dense_extent, float min_opacity,
                        float half_extent, float area_thr, uint8_t* out_a, uint8_t* out_b,
                        soar_stream_t stream);
/* Ascending list of the rows whose mask byte is non-zero (invert: zero) + their count (device). */
size_t soar_mask_index_workspace_bytes(int32_t N);
int soar_mask_to_index(const uint8_t* mask, int32_t N, int32_t invert, int32_t* index_out,
                       uint32_t* count_out, void* workspace, size_t workspace_bytes,
                       soar_stream_t stream);
/* soar_mask_to_index for callers that are not allowed to allocate device memory. It has no
 * counterpart in the reference (which builds these lists with torch boolean indexing and so always
 * allocates); it exists for hosts that reach this library from a place where cudaMalloc is off
 * limits: a plugin running inside another engine's memory pool with no allocator handed to it, a
 * real-time or signal-safe section, or a process whose device memory was fully budgeted up front.
 * Same inputs, outputs and results as soar_mask_to_index, but the per-CTA counts are kept in a
 * fixed array inside the library image instead of a caller workspace, so:
 *   - N is limited to SOAR_MASK_INDEX_STATIC_MAX_ROWS; more rows -> SOAR_E_WORKSPACE and a message
 *     pointing at soar_mask_to_index. The array is 16 KB per GPU, held from library load to exit.
 *   - the call itself makes no cudaMalloc / cudaFree / allocating library call (keep
 *     soar_prof_enable off: profiling creates events). The first call on each device looks up the
 *     array's device address (cudaGetSymbolAddress, with that device current); make that first
 *     call outside of any stream capture.
 *   - every call on a device uses the same array. Calls that follow each other on one stream are
 *     fine (nothing is carried over, nothing needs clearing); two calls that can run on the GPU at
 *     the same time (different streams, or different host threads on different streams) race on it
 *     and give wrong lists — serialise them, or use soar_mask_to_index with separate workspaces. */
#define SOAR_MASK_INDEX_STATIC_MAX_ROWS (1 << 24)
int soar_mask_to_index_static(const uint8_t* mask, int32_t N, int32_t invert, int32_t* index_out,
                              uint32_t* count_out, soar_stream_t stream);
#define SOAR_ROWS_MAX_TENSORS 32
typedef struct SoarRowTensor {
  float* data;               /* row-major [rows, width]; NULL source = zeros       */
  int32_t width;
  int32_t reserved;
} SoarRowTensor;
/* dst[t][dst_row0 + i, :] = src[t][index[i % n_index], :] for i < n_out and every tensor t, one
 * launch (index NULL: identity). Prune = gather of the kept rows; clone / split = gather of the
 * selected rows to the tail, with zero rows (NULL source) for the Adam moments (:905-977). */
int soar_rows_gather(const SoarRowTensor* src, const SoarRowTensor* dst, int32_t n_tensors,
                     const int32_t* index, int32_t n_index, int64_t n_out, int64_t dst_row0,
                     soar_stream_t stream);
/* densify_and_split's new centres and scales (:996-1010): for i < n_sel * n_rep, j = index[i % n_sel]:
 * out_xyz[i] = R(normalize(rotation[j])) . samples[i] + xyz[j];
 * out_scaling[i] = log(exp(scaling[j]) / (0.8 * n_rep)), z = -1e10 when surfel. */
int soar_densify_split_rows(const float* xyz, const float* scaling, const float* rotation,
                            const int32_t* index, int32_t n_sel, int32_t n_rep, const float* samples,
                            int32_t surfel, float* out_xyz, float* out_scaling, soar_stream_t stream);

/* ---- the reachable kernels of the reference's `cuda_utils` extension ----- */
/* gaussian2occgrid (TS/utils/cuda_utils.cu:779-872, pybind glue TS/utils/ext.cpp:201-255, caller
 * TS/geometry/surfel_base.py:1150-1195): grid[gx,gy,gz] is zeroed and every surfel adds
 * alpha = exp(-(dx^2+dy^2)/2) * opac to the cells its disc crosses. pos_min[3], pos_max[3], grid_len[1],
 * cutoff[1] are DEVICE scalars/vectors like the reference's tensors; grid_dim_host[3] is read on the
 * host (the reference calls .item() on it to size the grid, ext.cpp:214-216). pos[n,3], rot[n,4]
 * (w,x,y,z, used un-normalised), scale[n,3] (columns 0,1 used), opac[n]. error_flag (device u32) is
 * set to 1 when a Gaussian's sample lattice runs for more than 65536 in-bounds steps on an axis (a
 * grid_len that does not match the bounds: the reference would not return in reasonable time); such
 * Gaussians are skipped. One launch + two fills. */
int soar_occgrid_splat(int32_t n, const float* pos_min, const float* pos_max, const float* grid_len,
                       const float* cutoff, const int32_t* grid_dim_host, const float* pos,
                       const float* rot, const float* scale, const float* opac, float* grid,
                       uint32_t* error_flag, soar_stream_t stream);
/* contour_padding (TS/utils/cuda_utils.cu:226-300, glue TS/utils/ext.cpp:74-107, caller
 * TS/utils/image_utils.py:317): res[C,H,W] = 0, then every pixel outside mask[H,W] (bool bytes) takes
 * the mean over the masked pixels of its (2*size+1)^2 window. The reference's indexing is kept: only
 * plane 0 of res is written, from image[C*(min(15,C)-1) + pixel]. size == 0 returns the zeros. */
int soar_contour_padding(const float* image, const uint8_t* mask, int32_t C, int32_t H, int32_t W,
                         int32_t size, float* res, soar_stream_t stream);

/* ---- launch accounting / event timing (used by bench.py; no reference analogue:
 *      the reference has an unused CUDA-event Timer, TS/utils/cuda_utils.cu:18-42) --- */
/* Number of kernels this library has launched in this process. */
unsigned long long soar_launch_count(void);
/* on=1: start a fresh profile — every launch is bracketed by two events recorded
 * on its own stream. on=0: stop and drop the records. */
void soar_prof_enable(int on);
/* Synchronises the device, then fills counts[id] / total_ms[id] for id < max_ids
 * with the launches and summed device time seen since soar_prof_enable(1).
 * Returns the number of kernel ids. */
int soar_prof_summary(int max_ids, int* counts, float* total_ms);
const char* soar_kernel_name(int id);

/* ---- misc -------------------------------------------------------------------- */
const char* soar_last_error(void);
int soar_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SOAR_B200_H_INCLUDED */
