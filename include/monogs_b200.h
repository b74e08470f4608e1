/*
 * monogs_b200.h — C-ABI of the B200-native differentiable Gaussian-splatting rasterizer.
 *
 * Every entry point takes plain device pointers, explicit sizes, a caller-owned workspace and the
 * caller's CUDA stream (passed as void* so the header needs no CUDA include).  No torch type crosses
 * this boundary.  All functions return 0 on success or a negative MGS_E* code; the message of the
 * last failure on the calling thread is available from mgs_last_error_string().
 *
 * Paths below are relative to /root/reference/submodules/diff-gaussian-rasterization unless noted.
 * Matrices are 16 floats in the reference's memory order (element [4*col+row], auxiliary.h:58-76).
 */
#ifndef MONOGS_B200_H_
#define MONOGS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGS_OK 0
#define MGS_EINVAL (-1)    /* bad argument (null pointer, negative size, unsupported combination) */
#define MGS_ENOSPC (-2)    /* workspace too small for (P, W, H, R_cap) */
#define MGS_ECUDA (-3)     /* a CUDA runtime call failed; see mgs_last_error_string() */
#define MGS_EOVERFLOW (-4) /* num_rendered exceeded R_cap (only reported by the synchronous entry) */

/* gradient-request bits for mgs_raster_backward (the autograd wrapper sets them from needs_input_grad) */
#define MGS_GRAD_MEANS3D 1u
#define MGS_GRAD_COLOR 2u     /* dL_dsh or dL_dcolors_precomp */
#define MGS_GRAD_OPACITY 4u
#define MGS_GRAD_COV 8u       /* dL_dscales + dL_drotations, or dL_dcov3D_precomp */
#define MGS_GRAD_TAU 16u      /* camera pose tangent (rho, theta) */
#define MGS_GRAD_MEANS2D 32u  /* screen-space gradient sink used for densification statistics */
#define MGS_GRAD_ALL 63u

/* Per-view constants; mirrors GaussianRasterizationSettings (diff_gaussian_rasterization/__init__.py:392-405)
 * and the scalar arguments of _C.rasterize_gaussians (rasterize_points.h:18-39). */
typedef struct mgs_view {
  int32_t image_height, image_width;
  float tanfovx, tanfovy;
  float scale_modifier;
  int32_t sh_degree;      /* active degree D */
  int32_t sh_coeffs;      /* M = coefficients per channel in the shs tensor; 0 when colors_precomp is used */
  int32_t prefiltered;
  const float* bg;             /* device f32[3] */
  const float* viewmatrix;     /* device f32[16] */
  const float* projmatrix;     /* device f32[16] */
  const float* projmatrix_raw; /* device f32[16]; read only by backward (backward.cu:497-501) */
  const float* campos;         /* device; first three floats used (forward.cu:245) */
} mgs_view;

const char* mgs_last_error_string(void);
int mgs_version(void);
/* cudaStreamSynchronize(stream): the one host wait of the drop-in forward (the reference reads num_rendered back at
 * cuda_rasterizer/rasterizer_impl.cu:284) without a round trip through the caller's stream object. */
int mgs_stream_synchronize(void* stream);

/* Optional per-kernel timing with CUDA events on the launching stream (off by default; do not enable under
 * CUDA-graph capture).  mgs_profile_enable(1) clears and starts collecting, mgs_profile_report synchronises the
 * device and writes one line per kernel, "<name> <launches> <total_ms>\n", returning the number of lines. */
int mgs_profile_enable(int32_t on);
int mgs_profile_report(char* buf, size_t buf_bytes);

/* ---- rasterizer forward: replaces _C.rasterize_gaussians (ext.cpp:46, rasterize_points.cu:35-127,
 *      cuda_rasterizer/rasterizer_impl.cu:198-344).  The three growable torch arenas of the reference
 *      become one caller-owned workspace that forward fills and backward reads. ---- */

/* Bytes of workspace needed for P Gaussians, a W x H image and room for R_cap tile instances. */
size_t mgs_raster_workspace_bytes(int32_t P, int32_t W, int32_t H, int64_t R_cap);

/* Stage 1 (forward.cu:157-258 + the InclusiveSum at rasterizer_impl.cu:280): per-Gaussian projection,
 * conic, radius, tile rect, SH->RGB, the per-tile instance counts and their scan (= num_rendered).  Writes radii[P] and leaves
 * num_rendered in the workspace; if num_rendered_out is non-null it also receives the value through an
 * async copy on `stream` (device or pinned host pointer).  Optional inputs are null exactly where the
 * reference receives empty tensors (__init__.py:433-447). */
int mgs_raster_preprocess(const mgs_view* view, int32_t P,
                          const float* means3D, const float* opacities,
                          const float* shs, const float* colors_precomp,
                          const float* scales, const float* rotations, const float* cov3D_precomp,
                          int32_t* radii, void* workspace, size_t workspace_bytes,
                          int32_t* num_rendered_out, void* stream);

/* Stage 2 (rasterizer_impl.cu:286-341): binning — the reference's 64-bit (tile|depth) stable sort, done as a counting
 * sort on the tile half (tile ranges fall out of the scan) plus a per-tile sort of (depth bits, Gaussian id), which
 * yields the identical order — packed per-instance records and the alpha blend.  R_cap is the capacity the workspace
 * was sized for; if the device-side num_rendered exceeds it the overflow flag in the workspace is set, instances
 * beyond slot R_cap are dropped and the images are incomplete (re-run with a larger capacity).  Never synchronises.  out_* and n_touched are fully
 * written (n_touched is zeroed first), so the caller need not pre-fill them. */
int mgs_raster_render(const mgs_view* view, int32_t P, const float* colors_precomp, const int32_t* radii,
                      void* workspace, size_t workspace_bytes, int64_t R_cap,
                      float* out_color, float* out_depth, float* out_opacity, int32_t* n_touched,
                      void* stream);

/* Reads (num_rendered, overflow) back from a workspace; synchronises `stream`. */
int mgs_raster_query(const void* workspace, int32_t* num_rendered, int32_t* overflow, void* stream);

/* The whole forward of _C.rasterize_gaussians (rasterize_points.cu:35-127) in ONE call that never waits for the device:
 * stage 1 + stage 2 with the caller's GUESS of the capacity (the previous call's num_rendered plus headroom).  Where the
 * reference blocks on cudaMemcpy(num_rendered) in the middle of its forward (rasterizer_impl.cu:284), this call copies
 * status_out = i32[6] {num_rendered, overflow, r_used, n_large, overflow_sticky, max_rendered} (pinned host or device
 * memory) right after the scan and records ready_event (a cudaEvent_t, optional) behind that copy, BEFORE the emit / sort /
 * blend kernels are queued: the caller returns to its own work, waits for the event when it needs num_rendered (long fired
 * by then) and, in the rare case overflow != 0, grows the workspace (the per-Gaussian state sits at offsets that do not
 * depend on R_cap, so a copy of the first mgs_raster_workspace_bytes(P, W, H, 0) bytes carries it over) and calls
 * mgs_raster_render again.  Outputs and n_touched are fully written. */
int mgs_raster_forward(const mgs_view* view, int32_t P,
                       const float* means3D, const float* opacities,
                       const float* shs, const float* colors_precomp,
                       const float* scales, const float* rotations, const float* cov3D_precomp,
                       int32_t* radii, void* workspace, size_t workspace_bytes, int64_t R_cap,
                       float* out_color, float* out_depth, float* out_opacity, int32_t* n_touched,
                       int32_t* status_out, void* ready_event, void* stream);

/* Asynchronous copy of the first EIGHT header words of a workspace to pinned host or device memory: the six status
 * words (see status_out above), then [6] an internal ticket (zero between launches) and [7] max_tile_count, sticky like
 * max_rendered: the longest tile list (instances of one tile) any binning into this workspace saw — what a caller of
 * mgs_track_args.tile_segments compares with MGS_TILE_SEGMENT.  The two sticky words are only ever raised by the kernels: a loop that renders many views into one workspace
 * (mgs_track_iteration, mgs_map_view_*) polls them once in a while — no host wait inside the loop — and still learns
 * about ANY render that did not fit.  mgs_raster_status_reset clears the header (also required once before the first
 * use of a workspace whose sticky words are going to be read: the library never clears them itself). */
int mgs_raster_status_async(const void* workspace, int32_t* status_out /* i32[8] */, void* stream);
int mgs_raster_status_reset(void* workspace, void* stream);

/* ---- rasterizer backward: replaces _C.rasterize_gaussians_backward (ext.cpp:47,
 *      rasterize_points.cu:150-254, rasterizer_impl.cu:348-470, backward.cu).  Gradient outputs whose
 *      request bit is clear may be null.  All requested outputs are fully overwritten (no pre-zeroing
 *      needed).  dL_dtau_sum is f32[6] = sum over Gaussians of the reference's dL_dtau[P,6]
 *      (__init__.py:339-342: rho = [0:3], theta = [3:6]); dL_dtau_per_gaussian (f32[P,6]) is optional. ---- */
int mgs_raster_backward(const mgs_view* view, int32_t P, uint32_t grad_mask,
                        const float* means3D, const float* shs, const float* colors_precomp,
                        const float* scales, const float* rotations, const float* cov3D_precomp,
                        const int32_t* radii, const void* workspace, size_t workspace_bytes,
                        void* scratch, size_t scratch_bytes,
                        const float* dL_dout_color, const float* dL_dout_depth,
                        float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors_or_sh,
                        float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                        float* dL_dtau_sum, float* dL_dtau_per_gaussian, void* stream);
size_t mgs_raster_backward_scratch_bytes(int32_t P);

/* ---- pixel-sample backward: replaces the select_pixels = true branch of _C.rasterize_gaussians_backward
 *      (rasterize_points.cu:174-189,233; rasterizer_impl.cu:369,424; renderPixelSampleCUDA, backward.cu:592-797, launched
 *      with one block per listed pixel at backward.cu:1142).  Same outputs as mgs_raster_backward, but only the pixels in
 *      selected_pixel_indices (device i32[n_selected], linear index y * W + x) send gradient; a pixel listed k times
 *      sends it k times (the reference launches k blocks for it), indices outside [0, W*H) are ignored (its `inside`
 *      test).  n_selected = 0 yields zero gradients (the reference falls back to the dense kernel for an EMPTY list
 *      because it encodes "not selected" as size 0; callers wanting that use mgs_raster_backward).
 *      scratch: mgs_raster_backward_pixels_scratch_bytes(P, W, H). ---- */
int mgs_raster_backward_pixels(const mgs_view* view, int32_t P, uint32_t grad_mask,
                               const float* means3D, const float* shs, const float* colors_precomp,
                               const float* scales, const float* rotations, const float* cov3D_precomp,
                               const int32_t* radii, const void* workspace, size_t workspace_bytes,
                               void* scratch, size_t scratch_bytes,
                               const int32_t* selected_pixel_indices, int32_t n_selected,
                               const float* dL_dout_color, const float* dL_dout_depth,
                               float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors_or_sh,
                               float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                               float* dL_dtau_sum, float* dL_dtau_per_gaussian, void* stream);
size_t mgs_raster_backward_pixels_scratch_bytes(int32_t P, int32_t W, int32_t H);

/* ---- pose-only backward (tracking) and sketch-Jacobian backward.  Replaces the tracking use of
 *      _C.rasterize_gaussians_backward (only grad_theta / grad_rho are consumed there because the frontend's
 *      Gaussians are detached, utils/slam_frontend.py:1799-1800) and
 *      _C.rasterize_gaussians_backward_sketch_jacobian (ext.cpp:48, rasterize_points.cu:256-378,
 *      backward_sketch_jacobian.cu).  The screen-space-gradient -> pose-tangent map is linear per Gaussian, so
 *      each pixel accumulates its own 6-vector: tau_image[k,y,x] (f32[6,H,W], optional) is the pose tangent
 *      of that pixel's term dL_dout[:,y,x] . out[:,y,x]; dL_dtau_sum (f32[6], optional) is its sum over the
 *      image = sum over Gaussians of the reference's dL_dtau[P,6] (__init__.py:339-342). ---- */
int mgs_raster_backward_pose(const mgs_view* view, int32_t P,
                             const float* means3D, const float* shs, const float* colors_precomp,
                             const float* scales, const float* rotations, const float* cov3D_precomp,
                             const int32_t* radii, const void* workspace, size_t workspace_bytes,
                             void* scratch, size_t scratch_bytes,
                             const float* dL_dout_color, const float* dL_dout_depth,
                             float* tau_image, float* dL_dtau_sum, void* stream);
size_t mgs_raster_backward_pose_scratch_bytes(int32_t P, int32_t W, int32_t H);

/* Single-pass variant of the two calls above for the DROP-IN surface (what MonoGS's tracking actually calls: only theta / rho
 * require a gradient, utils/slam_frontend.py:1799-1800, and SH degree 0 so that no colour depends on the pose).  The pose
 * tangent of a pixel is tau = sum_ch g_ch (A_ch - out_ch E) (+ the depth terms) where E and A_ch are sums over the pixel's
 * splats that do NOT depend on the loss; mgs_raster_forward_pose is mgs_raster_forward that also accumulates them during
 * its one walk over the tile lists and leaves them in pose_state, and mgs_raster_backward_pose_deferred contracts them with
 * the dL/d out of whatever loss the caller evaluated in between — one pass over the pixels, no second walk over the lists,
 * no final_T / n_contrib round trip.  Same status_out / ready_event protocol as mgs_raster_forward; on overflow the caller
 * re-renders with mgs_raster_render and uses the two-pass mgs_raster_backward_pose. */
size_t mgs_raster_pose_state_bytes(int32_t P, int32_t W, int32_t H);
int mgs_raster_forward_pose(const mgs_view* view, int32_t P,
                            const float* means3D, const float* opacities,
                            const float* shs, const float* colors_precomp,
                            const float* scales, const float* rotations, const float* cov3D_precomp,
                            int32_t* radii, void* workspace, size_t workspace_bytes, int64_t R_cap,
                            float* out_color, float* out_depth, float* out_opacity, int32_t* n_touched,
                            void* pose_state, size_t pose_state_bytes,
                            int32_t* status_out, void* ready_event, void* stream);
int mgs_raster_backward_pose_deferred(int32_t W, int32_t H, const void* pose_state,
                                      const float* dL_dout_color, const float* dL_dout_depth /* may be null */,
                                      float* tau_image /* f32[6,H,W], optional */, float* dL_dtau_sum /* f32[6], optional */,
                                      void* stream);

/* Sketch rows from a tau image: sketch_dtau[j,i,:] = sum over pixels with sketch_indices[j,y,x] == i of
 * tau_image[:,y,x]; indices outside [0, sketch_dim) (the reference uses -1) are ignored.  This equals
 * sum over Gaussians of df_dtau[P,d,6] that the reference wrapper forms per stack j (__init__.py:216-253);
 * one tau image serves all n_stacks stacks because only the pixel->row routing differs between them. */
int mgs_sketch_reduce(const float* tau_image, int32_t H, int32_t W, int32_t n_stacks, int32_t sketch_dim,
                      const int32_t* sketch_indices /* i32[n_stacks,H,W] */,
                      float* sketch_dtau /* f32[n_stacks,sketch_dim,6] */, void* stream);

/* ---- markVisible: replaces _C.mark_visible (ext.cpp:49, rasterizer_impl.cu:54-66,141-153) ---- */
int mgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* ---- fused tracking / mapping loss (utils/slam_utils.py:83-113,188-253 and HuberLoss :58-75 of the
 *      MonoGS tree).  One pass computes the scalar loss and dL/d(image), dL/d(depth), dL/d(a), dL/d(b). ---- */
#define MGS_LOSS_TRACK_L1 0     /* get_loss_tracking_rgb / _rgbd: mean |opacity*(img_ab*m - gt*m)| */
#define MGS_LOSS_TRACK_HUBER 1  /* per-pixel residual -> Huber(delta) -> L2 norm (slam_frontend.py:506,596-598) */
#define MGS_LOSS_MAP_L1 2       /* get_loss_mapping_rgb / _rgbd */
typedef struct mgs_loss_args {
  int32_t mode, H, W, use_depth;
  float rgb_boundary_threshold; /* 0.01 */
  float alpha;                  /* rgbd mixing weight, 0.95 */
  float huber_delta;            /* 0.01 */
  float loss_scale;             /* multiplies the loss and all gradients (upstream grad of the scalar) */
  const float* image;    /* f32[3,H,W] rendered */
  const float* depth;    /* f32[1,H,W] rendered, null unless use_depth */
  const float* opacity;  /* f32[1,H,W] rendered */
  const float* gt_image; /* f32[3,H,W] */
  const float* gt_depth; /* f32[1,H,W] or null */
  const float* grad_mask;/* f32[1,H,W] (Camera.grad_mask) or null = all ones */
  const float* exposure_a; /* device f32[1] */
  const float* exposure_b; /* device f32[1] */
} mgs_loss_args;
/* out_loss: device f32[1]; out_dab: device f32[2] (dL/da, dL/db); grads may be null to skip backward. */
int mgs_loss_fwd_bwd(const mgs_loss_args* args, float* out_loss, float* out_dab,
                     float* dL_dimage, float* dL_ddepth, void* scratch, size_t scratch_bytes, void* stream);
size_t mgs_loss_scratch_bytes(int32_t H, int32_t W);

/* The same loss for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Inputs and
 * outputs are the caller's as above; the per-block partial sums live in a file-scope device array inside the library
 * (512 KB + 256 B, held by every GPU that loads the library until the process exits) whose device address the
 * library fetches once per device with cudaGetSymbolAddress.  Nothing on the call path calls cudaMalloc, cudaFree or
 * an allocating library routine.  Limit: H * W <= MGS_LOSS_STATIC_MAX_PIXELS, larger frames return MGS_ENOSPC with a
 * message (use mgs_loss_fwd_bwd with caller-owned scratch for those).  All calls on a device SHARE the array, which
 * keeps the previous call's contents: every slot a call reads it has written first, so calls ordered on one stream (or
 * otherwise serialised) are independent, but two calls that overlap on the GPU (different streams, no dependency
 * between them) race on it and return garbage.  Results are bit-identical to mgs_loss_fwd_bwd. */
#define MGS_LOSS_STATIC_MAX_PIXELS (2048 * 2048)
int mgs_loss_fwd_bwd_static(const mgs_loss_args* args, float* out_loss, float* out_dab,
                            float* dL_dimage, float* dL_ddepth, void* stream);

/* ---- fused Adam (torch.optim.Adam as configured at gaussian_splatting/scene/gaussian_model.py:285,
 *      utils/slam_frontend.py:453): multi-tensor, per-group lr, eps outside the bias-corrected sqrt. ---- */
typedef struct mgs_adam_group {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t numel; float lr;
  const float* lr_device;   /* optional device f32[1]: when non-null the learning rate is read from it at run time instead
                             * of `lr`, so a schedule (GaussianModel.update_learning_rate, gaussian_model.py:298-312: the xyz
                             * rate changes every iteration) works under CUDA-graph replay */
} mgs_adam_group;
int mgs_adam_step(const mgs_adam_group* groups /* host array */, int32_t n_groups, int32_t step /* t>=1 */,
                  float beta1, float beta2, float eps, void* stream);
/* Same update with the step number kept on the device: uses t = *step_counter + 1 and then increments the
 * counter, so an optimisation loop can be captured once in a CUDA graph and replayed for every iteration. */
int mgs_adam_step_device(const mgs_adam_group* groups, int32_t n_groups, int32_t* step_counter /* device i32[1] */,
                         float beta1, float beta2, float eps, void* stream);

/* ---- simple-knn: replaces simple_knn._C.distCUDA2 (submodules/simple-knn/spatial.cu:15-26,
 *      simple_knn.cu:46-222): mean squared distance to the three nearest other points. ---- */
size_t mgs_knn_workspace_bytes(int32_t P);
int mgs_knn_dist2(int32_t P, const float* points /* f32[P,3] */, float* mean_dist2 /* f32[P] */,
                  void* workspace, size_t workspace_bytes, void* stream);

/* ---- on-device pose update: replaces update_pose (utils/pose_utils.py:88-98, lietorch.SE3.exp).
 *      T_w2c (f32[4,4] row-major, in/out) <- Exp(rho, theta) @ T_w2c; rho and theta are zeroed; converged_out
 *      (optional) = |tau|^2 < threshold^2.  Optional outputs: viewmatrix = T^T (utils/camera_utils.py:94-96) and
 *      projmatrix = viewmatrix @ proj_raw (:98-104), so a tracking iteration needs no host round trip. ---- */
int mgs_pose_update(float* T_w2c, float* rho, float* theta, const float* proj_raw, float* viewmatrix_out,
                    float* projmatrix_out, int32_t* converged_out, float converged_threshold, void* stream);

/* ---- one whole tracking iteration, fused (the first-order inner loop of FrontEnd.tracking,
 *      utils/slam_frontend.py:455-630): render -> tracking loss (fused into the forward blend) -> pose-only backward
 *      -> Adam on (rho lr[0], theta lr[1], exposure_a lr[2], exposure_b lr[3]) -> T <- Exp(rho,theta) T.
 *      No host synchronisation; CUDA-graph capturable.  view.viewmatrix / view.projmatrix (and view.campos when it
 *      aliases the view matrix, as in MonoGS) are device buffers that the call UPDATES IN PLACE with the new pose.
 *      `workspace` and `scratch` must be ZERO-FILLED ONCE by their owner before the first iteration (and belong to this
 *      loop alone): tile counters, tickets and the tangent accumulator are left zero by every iteration for the next one
 *      (the sort kernel clears its tile's counter, the update kernel re-zeroes the accumulator), so an iteration is seven (four with tile_segments, below)
 *      kernel launches and no memset.  (MGS_SCAN_FOLD=1 / MGS_UPDATE_FOLD=1 in the environment run the tile scan / the
 *      update in the last CTA of the preceding kernel instead of as their own launches: measured slower, kept for comparison.)
 *      params = f32[8] (rho, theta, a, b); exp_avg / exp_avg_sq = f32[8]; step_counter = device i32[1]. ---- */
typedef struct mgs_track_args {
  mgs_view view;
  int32_t P;
  const float *means3D, *opacities, *shs, *scales, *rotations;
  int32_t loss_mode;   /* MGS_LOSS_TRACK_L1 or MGS_LOSS_TRACK_HUBER */
  int32_t use_depth;   /* rgbd variant of TRACK_L1 */
  int32_t fast_math;   /* 1: blend exponent via ex2.approx (rel. err ~4e-7) and n_touched NOT produced (left untouched);
                          0: expf as in the reference, n_touched written */
  float rgb_boundary_threshold, alpha, huber_delta;
  const float *gt_image, *gt_depth, *grad_mask;
  float* T_w2c;        /* f32[4,4] row-major, in/out */
  float *params, *exp_avg, *exp_avg_sq;
  int32_t* step_counter;
  float lr[4];
  float beta1, beta2, adam_eps, converged_threshold;
  float *out_color, *out_depth, *out_opacity;
  int32_t *radii, *n_touched;
  float *loss /* f32[1] */, *tau /* f32[6] */, *dab /* f32[2] */;
  int32_t* converged;  /* i32[1], optional */
  void* workspace; size_t workspace_bytes; int64_t R_cap;
  void* scratch; size_t scratch_bytes;
  /* optional.  lm_state: device f32[64], the tracker's cross-iteration record (layout below); with it the Huber
   * iteration also keeps the reference's best_viewpoint_params (utils/slam_frontend.py:523-531).  status_out: pinned host
   * i32[6], receives the workspace status words (see mgs_raster_status_async) as they stood when the call was enqueued,
   * i.e. of the previous iteration — an overflow watch that costs one 24-byte copy node and no host wait. */
  float* lm_state;
  int32_t* status_out;
  /* tile_segments = 1 (sh_degree 0 only): fixed-segment binning.  Every tile owns MGS_TILE_SEGMENT instance slots of the
   * workspace; the preprocess kernel scatters each (Gaussian, tile) instance into its tile's segment as it counts it, one
   * kernel sorts the segments, and the iteration is FOUR kernels (preprocess, tile_sort_segments, blend_track, track_update)
   * instead of seven — no tile scan and no emit pass between counting and sorting (the reference's duplicateWithKeys +
   * SortPairs + identifyTileRanges, rasterizer_impl.cu:288-321, as one scatter + one in-place sort).  Needs
   * R_cap >= tiles * MGS_TILE_SEGMENT.  Results are those of tile_segments = 0 as long as no tile holds more than
   * MGS_TILE_SEGMENT instances; a longer list loses its excess and raises the status words `overflow` / `overflow_sticky`
   * (mgs_raster_status_async), on which the caller falls back to tile_segments = 0. */
  int32_t tile_segments;
  int32_t reserved0;
} mgs_track_args;
#define MGS_TILE_SEGMENT 1024
size_t mgs_track_scratch_bytes(int32_t P, int32_t W, int32_t H);
int mgs_track_iteration(const mgs_track_args* args, void* stream);

/* ---- one SECOND-ORDER tracking iteration, fused (the sketched Gauss-Newton / Levenberg-Marquardt phase of
 *      FrontEnd.tracking, utils/slam_frontend.py:455-720 with in_second_order; gen_forward_sketch_args :269-338;
 *      ApplyExposure, utils/slam_utils.py:115-185; the `stack_dim` calls of _C.rasterize_gaussians_backward_sketch_jacobian,
 *      diff_gaussian_rasterization/__init__.py:216-253).  In the reference's order: apply the step left pending by the
 *      previous iteration (TempCamera.step + update_pose), render, scalar loss ||r||_1, adapt lambda, form the sketched
 *      Jacobian SJ[stack*sketch, 8] (rho, theta, exposure a, b) and residual Sf[stack*sketch] of
 *      f = rand_weights * sum_c Huber(r_c) / (npix / (stack*sketch)), solve lstsq([SJ; sqrt(lambda) I], -[Sf; 0]) and leave x
 *      pending.  One pass over the tile lists builds every pixel's row contribution (the per-pixel tangent is independent
 *      of the routing), one reduction routes them for all stacks: no [P, sketch, ...] buffers, no per-stack backward.
 *      Mono (rgb) objective and sh_degree 0 only, as the reference's own per-pixel rgbd loss raises NotImplementedError
 *      (slam_utils.py:221).  No host synchronisation; CUDA-graph capturable.
 *      lm_state layout (f32[64]): [0] lambda (set by the caller before the first iteration), [1] previous scalar loss,
 *      [2] have-previous flag, [3] step-pending flag, [4..11] pending step x (rho, theta, a, b), [12] best scalar loss,
 *      [13..28] best T_w2c, [29] best a, [30] best b, [31] converged (|x| < converged_threshold), [32] iterations done,
 *      [33] scalar loss of the last iteration, [34] |x|, [35] have-best flag.  Zero it (except [0]) to start a frame. ---- */
typedef struct mgs_track2_args {
  mgs_track_args t;              /* loss_mode / use_depth / lr / Adam moments / tau / dab are not used; t.loss receives ||r||_1 */
  int32_t n_stacks, sketch_dim;  /* stack_dim, sketch_dim (<= 64) */
  const int32_t* sketch_indices; /* device i32[n_stacks,H,W]: row of the pixel in that stack, -1 = not in the stack */
  const float* rand_weights;     /* device f32[H,W], +-1 */
  float* rows;                   /* out, device f32[n_stacks*sketch_dim, 9]: columns 0..7 = SJ, column 8 = Sf */
  float* lm_state;               /* device f32[64]; null = t.lm_state */
  float min_lambda, max_lambda, increase_factor, decrease_factor, converged_threshold;
} mgs_track2_args;
int mgs_track_second_order(const mgs_track2_args* args, void* stream);

/* ---- keyframe-window mapping iteration on the RAW Gaussian parameters (SURVEY.md §8 rows f1 + e).
 *      Replaces, per iteration of BackEnd.map (utils/slam_backend.py:157-333): the per-render parameter activations and
 *      their autograd chain (gaussian_splatting/scene/gaussian_model.py:77-97, gaussian_renderer/__init__.py:79-114:
 *      sigmoid(_opacity), exp(_scaling), normalize(_rotation), cat(f_dc, f_rest)), render + get_loss_mapping
 *      (utils/slam_utils.py:224-253) + backward for every view, the isotropic regulariser (:244-246) and the
 *      densification statistics (:292-299, gaussian_model.py:693-697).  The activations run ONCE per iteration
 *      (mgs_map_begin), every view accumulates activated-space gradients in place (mgs_map_view, no host
 *      synchronisation), and mgs_map_end applies the activation chain rule once and writes the raw-parameter gradients
 *      — the arena NCCL all-reduces and the fused Adam consumes. ---- */
typedef struct mgs_gaussians {
  int32_t P;
  int32_t sh_coeffs;        /* M: coefficients per channel stored (1 + rest) */
  int32_t sh_degree;        /* active degree */
  int32_t isotropic;        /* 1: scaling_raw is [P,1] and is repeated to three axes (gaussian_renderer/__init__.py:92-95) */
  const float* xyz;          /* f32[P,3] */
  const float* f_dc;         /* f32[P,1,3] */
  const float* f_rest;       /* f32[P,M-1,3], null when M == 1 */
  const float* opacity_raw;  /* f32[P,1], pre-sigmoid */
  const float* scaling_raw;  /* f32[P,3] (or [P,1]), pre-exp */
  const float* rotation_raw; /* f32[P,4], un-normalised (w,x,y,z) */
} mgs_gaussians;

typedef struct mgs_map_args {
  mgs_gaussians g;
  /* raw-parameter gradients, same shapes as the parameters.  grad_xyz, grad_f_dc and grad_f_rest are accumulated in place by the
   * views (zeroed by mgs_map_begin); the others are written by mgs_map_end. */
  float *grad_xyz, *grad_f_dc, *grad_f_rest, *grad_opacity, *grad_scaling, *grad_rotation;
  /* densification statistics updated in place by every view with update_stats != 0; each may be null */
  float *xyz_gradient_accum, *denom, *max_radii2D;   /* f32[P] each */
  float* loss_total;        /* device f32[1]: sum of the view losses (+ regulariser), zeroed by mgs_map_begin */
  void* scratch; size_t scratch_bytes;   /* mgs_map_scratch_bytes(P, M, W_max, H_max) */
} mgs_map_args;

typedef struct mgs_map_view {
  mgs_view view;
  const float* gt_image;    /* f32[3,H,W] */
  const float* gt_depth;    /* f32[1,H,W], null = monocular (get_loss_mapping_rgb) */
  const float* exposure_a;  /* device f32[1] or null (initialization: image_ab = image) */
  const float* exposure_b;
  float rgb_boundary_threshold, alpha;
  int32_t update_stats;
  int32_t fast_math;        /* 1: ex2.approx exponent in this view's forward and backward blend (rel. error ~4e-7) */
  int32_t slot;             /* 0..3: which copy of the per-view loss-gradient images in args->scratch this view uses, so
                             * that the forward of one view may run beside the backward of another (different streams,
                             * different workspaces, different slots) */
  int32_t concurrent_backward; /* 1: this view's backward accumulates gradients and statistics with atomics, so that
                             * backward calls of different views (different slots) may run on different streams */
  float *out_color, *out_depth, *out_opacity;   /* f32[3,H,W], f32[1,H,W], f32[1,H,W] */
  int32_t *radii, *n_touched;                   /* i32[P] of THIS view (visibility filter, occ_aware_visibility) */
  float* tau;               /* f32[6]: pose tangent (rho, theta) of this view's loss, optional */
  float* dab;               /* f32[2]: dL/d(exposure_a, exposure_b), optional */
  float* loss;              /* f32[1]: this view's loss, optional */
  void* workspace; size_t workspace_bytes; int64_t R_cap;   /* zero-filled once by its owner before the first view
                             * rendered into it, and used by mgs_map_view_* only (tile counters and the scan ticket are left
                             * zero by every view for the next one) */
  int32_t tile_segments;    /* 1: fixed-segment binning as in mgs_track_args.tile_segments (R_cap >= tiles * MGS_TILE_SEGMENT;
                             * a tile list beyond a segment raises the workspace's overflow words): per view preprocess,
                             * tile_sort_segments, blend, ... instead of preprocess + scan, emit, two sort kernels, blend, ... */
  int32_t reserved0;
} mgs_map_view;

size_t mgs_map_scratch_bytes(int32_t P, int32_t sh_coeffs, int32_t W_max, int32_t H_max);
int mgs_map_begin(const mgs_map_args* args, void* stream);
int mgs_map_view_fwd_bwd(const mgs_map_args* args, const mgs_map_view* view, void* stream);
/* The two halves of mgs_map_view_fwd_bwd.  forward: preprocess, binning, blend + mapping loss (writes the view's
 * workspace, radii, n_touched, loss, dab and slot `view->slot` of the loss-gradient images).  backward: backward blend +
 * per-Gaussian backward accumulating into args' gradients and statistics, tau.  All backward calls of an iteration
  * must be ordered among themselves (one stream) unless every view sets concurrent_backward; a forward may overlap the backward or the forward of another view
 * (different streams, workspaces, slots and output images). */
int mgs_map_view_forward(const mgs_map_args* args, const mgs_map_view* view, void* stream);
int mgs_map_view_backward(const mgs_map_args* args, const mgs_map_view* view, void* stream);
/* add_regulariser: 1 on exactly one rank of a data-parallel group so the all-reduced sum counts the isotropic term
 * weight * mean|s - mean(s)| once (weight = 10 in the reference). */
int mgs_map_end(const mgs_map_args* args, int32_t add_regulariser, float regulariser_weight, void* stream);

/* Window poses of a mapping iteration: keyframe_optimizers.step() followed by update_pose(viewpoint) for the keyframes
 * being optimised (utils/slam_backend.py:325-332; optimiser built at :437-489 — torch.optim.Adam with its default eps,
 * cam_rot_delta / cam_trans_delta at half the tracking rates, exposure a / b at 0.01).  One call handles all the views a
 * rank owns (n <= 16): Adam on (rho, theta, a, b) with the view's tau / dab of mgs_map_view_backward, T <- Exp(rho, theta) T,
 * rho = theta = 0, and the view's viewmatrix / projmatrix rewritten in place so that the next iteration renders from the
 * new pose without a host round trip.  gather_out (optional, f32[20] = T | a | b | 0 | 0) is the owner's contribution to
 * the pose all-gather of a sharded window: it points into the trailing sum-only segment of the gradient arena, the other
 * ranks contribute zeros there, and the one exchange of the iteration (NCCL all-reduce or mgs_peer_reduce_adam) leaves
 * every pose on every rank.  frozen = 1 (the first keyframe, uid 0): no step, matrices and gather_out still written. */
typedef struct mgs_pose_slot {
  float* T_w2c;             /* f32[4,4] row-major, in/out */
  float* params;            /* f32[8]: rho, theta, exposure a, b; in/out */
  float* exp_avg; float* exp_avg_sq;   /* f32[8] each */
  int32_t* step;            /* device i32[1]: Adam step count of this view's optimiser */
  const float* tau;         /* f32[6]: d loss / d (rho, theta) of this iteration */
  const float* dab;         /* f32[2]: d loss / d (a, b) */
  float* viewmatrix; float* projmatrix;   /* f32[16] each, optional */
  float* gather_out;        /* f32[20], optional */
  float lr[4];              /* rho, theta, a, b */
  int32_t frozen;
} mgs_pose_slot;
int mgs_window_pose_step(const mgs_pose_slot* slots /* host array */, int32_t n, const float* proj_raw /* device f32[16] */,
                         float beta1, float beta2, float eps, void* stream);

/* ---- peer-memory exchange of the data-parallel mapping step.  The reference has no multi-GPU path
 *      (BackEnd.map, utils/slam_backend.py:157-333, renders the whole window on one GPU and calls
 *      torch.optim.Adam.step, gaussian_model.py:285); in the sharded window every rank holds gradients of its own views
 *      and the step that follows is "sum over ranks, then Adam".  mgs_peer_reduce_adam does both in one kernel over
 *      NVLink peer mappings: rank r owns slice r of the flat arena, reads that slice of EVERY rank's gradient arena
 *      (summed in rank order), applies Adam to it (moments of a slice live on its owner only) and stores the new
 *      parameters into EVERY rank's parameter arena; the trailing loss element is summed into every rank's loss_out.
 *      A second one-warp kernel holds the stream until all ranks have finished, so the caller's next launch may reuse
 *      the gradient arena and read the parameters.  No host synchronisation, no NCCL call.
 *      Blocks come from mgs_peer_alloc (cudaMalloc, zero filled) and are mapped into the other ranks' address space
 *      with mgs_peer_export / mgs_peer_open (CUDA IPC; the 64-byte handles travel by any host channel). ---- */
#define MGS_PEER_MAX 8
#define MGS_PEER_MAX_SEGMENTS 16
#define MGS_PEER_HANDLE_BYTES 64
int mgs_peer_alloc(size_t bytes, void** out_ptr);
int mgs_peer_free(void* ptr);
int mgs_peer_export(const void* ptr, uint8_t* handle /* host u8[MGS_PEER_HANDLE_BYTES] */);
int mgs_peer_open(const uint8_t* handle, void** out_ptr);
int mgs_peer_close(void* ptr);
size_t mgs_peer_flags_bytes(void);   /* size of one rank's flag block (zero filled before first use) */
typedef struct mgs_peer_exchange {
  int32_t world, rank;                      /* 1..MGS_PEER_MAX ranks; entries [0, world) of the tables are used */
  float* param_arena[MGS_PEER_MAX];         /* f32[numel] of every rank, as mapped in THIS process; [rank] is local */
  const float* grad_arena[MGS_PEER_MAX];    /* f32[numel], same layout */
  float* loss_out[MGS_PEER_MAX];            /* per rank, as many floats as the trailing sum-only segment(s) hold (>= 4):
                                             * receives their sums — the iteration's loss, and whatever else rides along
                                             * (the window poses each owner contributes, zeros elsewhere) */
  uint32_t* flags[MGS_PEER_MAX];            /* mgs_peer_flags_bytes() per rank */
  float* exp_avg; float* exp_avg_sq;        /* local f32[numel]; only this rank's slice is read and written */
  int64_t numel;                            /* arena length in floats, multiple of 4 */
  int32_t n_segments;                       /* consecutive segments covering [0, numel): one per parameter tensor */
  int64_t segment_end[MGS_PEER_MAX_SEGMENTS];   /* exclusive ends in floats, multiples of 4 */
  float segment_lr[MGS_PEER_MAX_SEGMENTS];      /* learning rate of the segment; < 0: sum only (trailing segments only) */
  /* Optional NVSwitch multicast (NVLS) mappings of the same memory — one address that reaches the arena of EVERY rank
   * (cuMulticast* objects; torch.distributed._symmetric_memory provides them).  All three or none.  When given, the
   * gradients are read with multimem.ld_reduce (summed inside the switch, association order unspecified) and the
   * parameters / loss are written with multimem.st; the unicast tables above are still used for the flags, the own
   * parameters and nothing else. */
  const float* grad_multicast; float* param_multicast; float* loss_multicast;
  uint32_t timeout_ms;   /* how long a rank waits on the device for its peers before it gives up, records the epoch in
                          * its error word (mgs_peer_error) and goes on with whatever it has; 0 = 60 000 */
} mgs_peer_exchange;
/* step: Adam step number t >= 1 (identical on all ranks); epoch: strictly increasing per call on the same blocks,
 * identical on all ranks, first call 1.  Slice r = float4 units [r*ceil(numel/4/world), ...). */
int mgs_peer_reduce_adam(const mgs_peer_exchange* x, int32_t step, uint32_t epoch, float beta1, float beta2, float eps,
                         void* stream);
/* Synchronises `stream` and returns the epoch at which a wait of this rank timed out (0 = none). */
int mgs_peer_error(const uint32_t* flags, uint32_t* out_epoch, void* stream);

/* ---- densify / clone / split / prune of the parameter arena together with its Adam moments.
 *      Replaces GaussianModel.densify_and_prune / densify_and_clone / densify_and_split / prune_points and the optimizer
 *      surgery of cat_tensors_to_optimizer / _prune_optimizer (gaussian_splatting/scene/gaussian_model.py:470-691).
 *      mgs_densify_plan decides per source Gaussian and scans; the host reads the four totals (the new size is needed
 *      to allocate), then mgs_densify_apply writes the final arena
 *          [kept originals | clones | children rep 0 | .. | children rep n_split-1]        (each in source order)
 *      which is the order the reference's cat / boolean-mask sequence produces.  New points get zero Adam moments. ---- */
typedef struct mgs_densify_cfg {
  int32_t mode;            /* 0: densify_and_prune(max_grad, min_opacity, extent, max_screen_size); 1: prune_points(mask) */
  int32_t n_split;         /* N of densify_and_split (2) */
  float grad_threshold;    /* max_grad */
  float min_opacity;
  float extent;
  float percent_dense;     /* 0.01 */
  float max_screen_size;   /* 0 = None: the world-size prune (max scale > 0.1 extent) is applied only when non-zero */
} mgs_densify_cfg;

typedef struct mgs_gaussian_state {   /* arrays in the order xyz, f_dc, f_rest, opacity, scaling, rotation */
  float* param[6];
  float* exp_avg[6];       /* Adam moments; entries may be null (no optimizer state yet) */
  float* exp_avg_sq[6];
  int32_t* unique_kfIDs;   /* i32[P], optional */
  int32_t* n_obs;          /* i32[P], optional */
} mgs_gaussian_state;

size_t mgs_densify_scratch_bytes(int32_t P);
/* totals_out (device or pinned host i32[4], written by an async copy on `stream`):
 * [0] kept originals, [1] clones, [2] children PER REPETITION, [3] split-selected Gaussians (rows of noise per repetition).
 * New size = totals[0] + totals[1] + n_split * totals[2].  xyz_gradient_accum may be null (no gradient statistics). */
int mgs_densify_plan(const mgs_gaussians* g, const float* xyz_gradient_accum, const float* denom, const uint8_t* prune_mask,
                     const mgs_densify_cfg* cfg, void* scratch, size_t scratch_bytes, int32_t* totals_out, void* stream);
/* noise: f32[n_split * totals[3], 3] standard-normal draws, row r * totals[3] + k for repetition r of the k-th selected
 * Gaussian (the layout of `samples` at gaussian_model.py:612-614); may be null when totals[3] == 0 or mode == 1. */
int mgs_densify_apply(const mgs_gaussians* g, const mgs_gaussian_state* src, const mgs_gaussian_state* dst,
                      const mgs_densify_cfg* cfg, const float* noise, const void* scratch, size_t scratch_bytes, void* stream);

/* ---- per-frame image preprocessing and covisibility bookkeeping ---- */
/* Exact lower median (torch.median) of n floats by radix select on the float bits; NaNs are skipped; device result. */
size_t mgs_median_scratch_bytes(void);
int mgs_median(const float* x, int64_t n, float* out /* device f32[1] */, void* scratch, size_t scratch_bytes, void* stream);
/* get_median_depth (utils/slam_utils.py:286-297): median of depth over depth > 0 [and opacity > 0.95] [and mask != 0];
 * opacity / mask / n_valid_out may be null.  NaN when nothing is valid. */
int mgs_median_depth(const float* depth, const float* opacity, const float* mask, int64_t n, float* out, int32_t* n_valid_out,
                     void* scratch, size_t scratch_bytes, void* stream);
/* sqrt(gv^2 + gh^2) of the Scharr gradients of mean_c(image), zero where the 3x3 neighbourhood of |gray| > eps is not
 * full (image_gradient + image_gradient_mask, utils/slam_utils.py:7-41; Camera.compute_grad_mask, camera_utils.py:110-118). */
int mgs_frame_grad_intensity(const float* image /* f32[C,H,W] */, int32_t C, int32_t H, int32_t W, float eps,
                             float* intensity /* f32[H,W] */, void* stream);
/* Camera.compute_grad_mask (utils/camera_utils.py:110-142): grad_mask = intensity > median * edge_threshold (global median;
 * patch_median != 0: per 32x32 patch as for Replica), rgb_pixel_mask = (sum_c image > thr) * grad_mask,
 * rgb_pixel_mask_mapping = (sum_c image > thr).  Masks are f32[1,H,W] of 0/1; the last two may be null.
 * scratch: mgs_median_scratch_bytes() + 256 bytes. */
int mgs_frame_masks(const float* image /* f32[3,H,W] */, int32_t H, int32_t W, float edge_threshold, float rgb_boundary_threshold,
                    int32_t patch_median, float* intensity_scratch /* f32[H,W] */, float* grad_mask, float* rgb_pixel_mask,
                    float* rgb_pixel_mask_mapping, void* scratch, size_t scratch_bytes, void* stream);
/* bits[w] bit b = n_touched[32 w + b] > 0 (occ_aware_visibility, utils/slam_backend.py:251-255); u32[(P+31)/32]. */
int mgs_visibility_pack(int32_t P, const int32_t* n_touched, uint32_t* bits, void* stream);
/* out i32[K+1,3]: row k < K = (|cur & kf_k|, |cur | kf_k|, |kf_k|), row K = (|cur|, 0, 0): the counts behind is_keyframe /
 * add_to_window / the mapping overlap check (utils/slam_frontend.py:1713-1741, 1918-1932).  kf_bits is u32[K, words]. */
int mgs_covisibility(int32_t P, const uint32_t* cur_bits, int32_t K, const uint32_t* kf_bits, int32_t* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MONOGS_B200_H_ */
