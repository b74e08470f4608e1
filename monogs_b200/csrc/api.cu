// extern "C" surface of libmonogs_b200 (include/monogs_b200.h): argument validation, workspace carving and
// kernel sequencing.  Nothing here synchronises except mgs_raster_query.
#include "launchers.h"
#include <cstring>

using namespace mgs;

namespace mgs {
void prof_enable(bool on);
int prof_report(char* buf, size_t n);
int launch_backward_full(const BackwardArgs& a, cudaStream_t s);
int launch_backward_pose(const BackwardArgs& a, cudaStream_t s);
size_t backward_pose_scratch_bytes(int P, int W, int H);
int launch_sketch_reduce(const float* rows, int H, int W, int n_stacks, int d, const int32_t* sketch_indices,
                         float* out, cudaStream_t s, int nv);
}

static int check_view(const mgs_view* v, bool need_raw) {
  if (!v) { set_error("view is null"); return MGS_EINVAL; }
  if (v->image_height <= 0 || v->image_width <= 0) { set_error("image size must be positive"); return MGS_EINVAL; }
  if (!v->bg || !v->viewmatrix || !v->projmatrix) { set_error("bg/viewmatrix/projmatrix must be non-null"); return MGS_EINVAL; }
  if (need_raw && !v->projmatrix_raw) { set_error("projmatrix_raw must be non-null for backward"); return MGS_EINVAL; }
  if (v->sh_degree < 0 || v->sh_degree > 3) { set_error("sh_degree must be in [0,3]"); return MGS_EINVAL; }
  return MGS_OK;
}

static int check_options(const mgs_view* v, const float* shs, const float* colors_precomp, const float* scales,
                         const float* rotations, const float* cov3D_precomp) {
  if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("provide exactly one of shs / colors_precomp"); return MGS_EINVAL; }
  if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) {
    set_error("provide exactly one of (scales, rotations) / cov3D_precomp"); return MGS_EINVAL;
  }
  if (shs) {
    const int need = (v->sh_degree + 1) * (v->sh_degree + 1);
    if (v->sh_coeffs < need) { set_error("sh_coeffs=%d < (sh_degree+1)^2=%d", v->sh_coeffs, need); return MGS_EINVAL; }
    if (!v->campos && v->sh_degree > 0) { set_error("campos required for sh_degree > 0"); return MGS_EINVAL; }
  }
  return MGS_OK;
}

extern "C" {

const char* mgs_last_error_string(void) { return mgs::get_error(); }
int mgs_version(void) { return 100; }

int mgs_stream_synchronize(void* stream) {
  MGS_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  return MGS_OK;
}
int mgs_profile_enable(int32_t on) { prof_enable(on != 0); return MGS_OK; }
int mgs_profile_report(char* buf, size_t buf_bytes) { if (!buf) { set_error("null buffer"); return MGS_EINVAL; } return prof_report(buf, buf_bytes); }

size_t mgs_raster_workspace_bytes(int32_t P, int32_t W, int32_t H, int64_t R_cap) {
  WsLayout L;
  if (ws_layout(P, W, H, R_cap, &L) != MGS_OK) return 0;
  return L.total;
}

int mgs_raster_preprocess(const mgs_view* view, int32_t P, const float* means3D, const float* opacities,
                          const float* shs, const float* colors_precomp, const float* scales,
                          const float* rotations, const float* cov3D_precomp, int32_t* radii, void* workspace,
                          size_t workspace_bytes, int32_t* num_rendered_out, void* stream) {
  int rc = check_view(view, false);
  if (rc) return rc;
  if (P < 0) { set_error("P < 0"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  if (!workspace) { set_error("workspace is null"); return MGS_EINVAL; }
  WsLayout L;
  if ((rc = ws_layout(P, view->image_width, view->image_height, 0, &L))) return rc;
  if (workspace_bytes < L.total) { set_error("workspace too small: %zu < %zu", workspace_bytes, L.total); return MGS_ENOSPC; }
  WsHeader* hdr = ws_ptr<WsHeader>(workspace, L.header);
  if (P == 0) {
    MGS_CUDA_TRY(cudaMemsetAsync(hdr, 0, sizeof(WsHeader), s));
    if (num_rendered_out) MGS_CUDA_TRY(cudaMemsetAsync(num_rendered_out, 0, sizeof(int32_t), s));
    return MGS_OK;
  }
  if (!means3D || !opacities || !radii) { set_error("means3D/opacities/radii must be non-null"); return MGS_EINVAL; }
  if ((rc = check_options(view, shs, colors_precomp, scales, rotations, cov3D_precomp))) return rc;
  const ViewConst v = make_view_const(view);
  if ((rc = launch_preprocess(v, P, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                              workspace, L, s)))
    return rc;
  if ((rc = launch_tile_scan(v, workspace, L, 0xFFFFFFFFLL, s))) return rc;   // header.num_rendered = sum of the tile counts
  if (num_rendered_out) MGS_CUDA_TRY(cudaMemcpyAsync(num_rendered_out, &hdr->num_rendered, sizeof(uint32_t), cudaMemcpyDefault, s));
  return MGS_OK;
}

int mgs_raster_render(const mgs_view* view, int32_t P, const float* colors_precomp, const int32_t* radii,
                      void* workspace, size_t workspace_bytes, int64_t R_cap, float* out_color, float* out_depth, float* out_opacity,
                      int32_t* n_touched, void* stream) {
  int rc = check_view(view, false);
  if (rc) return rc;
  if (P < 0 || R_cap < 0 || R_cap > 0x7fffffffLL) { set_error("bad P or R_cap"); return MGS_EINVAL; }
  if (!workspace || !out_color || !out_depth || !out_opacity) { set_error("null output/workspace"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  const int W = view->image_width, H = view->image_height;
  if (P == 0) {  // rasterize_points.cu:90 — zero-filled outputs, no kernels
    MGS_CUDA_TRY(cudaMemsetAsync(out_color, 0, (size_t)3 * W * H * 4, s));
    MGS_CUDA_TRY(cudaMemsetAsync(out_depth, 0, (size_t)W * H * 4, s));
    MGS_CUDA_TRY(cudaMemsetAsync(out_opacity, 0, (size_t)W * H * 4, s));
    return MGS_OK;
  }
  if (!n_touched || !radii) { set_error("n_touched/radii is null"); return MGS_EINVAL; }
  WsLayout L;
  if ((rc = ws_layout(P, W, H, R_cap, &L))) return rc;
  if (workspace_bytes < L.total) { set_error("workspace too small: %zu < %zu", workspace_bytes, L.total); return MGS_ENOSPC; }
  const ViewConst v = make_view_const(view);
  MGS_CUDA_TRY(cudaMemsetAsync(n_touched, 0, (size_t)P * sizeof(int32_t), s));
  if ((rc = launch_binning(v, P, colors_precomp, radii, workspace, L, R_cap, s))) return rc;
  return launch_blend_forward(v, workspace, L, out_color, out_depth, out_opacity, n_touched, s);
}

int mgs_raster_query(const void* workspace, int32_t* num_rendered, int32_t* overflow, void* stream) {
  if (!workspace) { set_error("workspace is null"); return MGS_EINVAL; }
  WsHeader h;
  cudaStream_t s = (cudaStream_t)stream;
  MGS_CUDA_TRY(cudaMemcpyAsync(&h, workspace, sizeof(uint32_t) * 4, cudaMemcpyDeviceToHost, s));
  MGS_CUDA_TRY(cudaStreamSynchronize(s));
  if (num_rendered) *num_rendered = (int32_t)h.num_rendered;
  if (overflow) *overflow = (int32_t)h.overflow;
  return MGS_OK;
}

int mgs_raster_status_async(const void* workspace, int32_t* status_out, void* stream) {
  if (!workspace || !status_out) { set_error("raster_status: null argument"); return MGS_EINVAL; }
  MGS_CUDA_TRY(cudaMemcpyAsync(status_out, workspace, 8 * sizeof(uint32_t), cudaMemcpyDefault, (cudaStream_t)stream));
  return MGS_OK;
}

int mgs_raster_status_reset(void* workspace, void* stream) {
  if (!workspace) { set_error("raster_status_reset: workspace is null"); return MGS_EINVAL; }
  MGS_CUDA_TRY(cudaMemsetAsync(workspace, 0, sizeof(WsHeader), (cudaStream_t)stream));
  return MGS_OK;
}

int mgs_raster_forward(const mgs_view* view, int32_t P, const float* means3D, const float* opacities,
                       const float* shs, const float* colors_precomp, const float* scales, const float* rotations,
                       const float* cov3D_precomp, int32_t* radii, void* workspace, size_t workspace_bytes, int64_t R_cap,
                       float* out_color, float* out_depth, float* out_opacity, int32_t* n_touched,
                       int32_t* status_out, void* ready_event, void* stream) {
  int rc = check_view(view, false);
  if (rc) return rc;
  if (P < 0 || R_cap < 0 || R_cap > 0x7fffffffLL) { set_error("bad P or R_cap"); return MGS_EINVAL; }
  if (!workspace || !out_color || !out_depth || !out_opacity) { set_error("null output/workspace"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  const int W = view->image_width, H = view->image_height;
  WsLayout L;
  if ((rc = ws_layout(P, W, H, R_cap, &L))) return rc;
  if (workspace_bytes < L.total) { set_error("workspace too small: %zu < %zu", workspace_bytes, L.total); return MGS_ENOSPC; }
  if (P == 0) {  // rasterize_points.cu:90 — zero-filled outputs, no kernels
    MGS_CUDA_TRY(cudaMemsetAsync(workspace, 0, sizeof(WsHeader), s));
    MGS_CUDA_TRY(cudaMemsetAsync(out_color, 0, (size_t)3 * W * H * 4, s));
    MGS_CUDA_TRY(cudaMemsetAsync(out_depth, 0, (size_t)W * H * 4, s));
    MGS_CUDA_TRY(cudaMemsetAsync(out_opacity, 0, (size_t)W * H * 4, s));
    if (status_out) MGS_CUDA_TRY(cudaMemcpyAsync(status_out, workspace, 6 * sizeof(uint32_t), cudaMemcpyDefault, s));
    if (ready_event) MGS_CUDA_TRY(cudaEventRecord((cudaEvent_t)ready_event, s));
    return MGS_OK;
  }
  if (!means3D || !opacities || !radii || !n_touched) { set_error("means3D/opacities/radii/n_touched must be non-null"); return MGS_EINVAL; }
  if ((rc = check_options(view, shs, colors_precomp, scales, rotations, cov3D_precomp))) return rc;
  const ViewConst v = make_view_const(view);
  if ((rc = launch_preprocess(v, P, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                              workspace, L, s)))
    return rc;
  MGS_CUDA_TRY(cudaMemsetAsync(n_touched, 0, (size_t)P * sizeof(int32_t), s));
  if ((rc = launch_tile_scan(v, workspace, L, R_cap, s))) return rc;
  // the verdict of the scan goes out before the heavy kernels are queued behind it
  if (status_out) MGS_CUDA_TRY(cudaMemcpyAsync(status_out, workspace, 6 * sizeof(uint32_t), cudaMemcpyDefault, s));
  if (ready_event) MGS_CUDA_TRY(cudaEventRecord((cudaEvent_t)ready_event, s));
  if ((rc = launch_binning(v, P, colors_precomp, radii, workspace, L, R_cap, s, false, /*scan_done=*/true))) return rc;
  return launch_blend_forward(v, workspace, L, out_color, out_depth, out_opacity, n_touched, s);
}

size_t mgs_raster_pose_state_bytes(int32_t P, int32_t W, int32_t H) {
  if (P < 0 || W <= 0 || H <= 0) return 0;
  return pose_deferred_bytes(W, H) + ((size_t)(P > 0 ? P : 1) * 32 * sizeof(float) + 255) / 256 * 256;
}

int mgs_raster_forward_pose(const mgs_view* view, int32_t P, const float* means3D, const float* opacities,
                            const float* shs, const float* colors_precomp, const float* scales, const float* rotations,
                            const float* cov3D_precomp, int32_t* radii, void* workspace, size_t workspace_bytes, int64_t R_cap,
                            float* out_color, float* out_depth, float* out_opacity, int32_t* n_touched,
                            void* pose_state, size_t pose_state_bytes, int32_t* status_out, void* ready_event, void* stream) {
  int rc = check_view(view, true);
  if (rc) return rc;
  if (P <= 0 || R_cap <= 0 || R_cap > 0x7fffffffLL) { set_error("forward_pose: bad P or R_cap"); return MGS_EINVAL; }
  if (view->sh_degree != 0) { set_error("forward_pose: sh_degree must be 0 (the colour of a splat must not depend on the pose)"); return MGS_EINVAL; }
  if (!workspace || !out_color || !out_depth || !out_opacity || !means3D || !opacities || !radii || !n_touched || !pose_state) {
    set_error("forward_pose: null argument"); return MGS_EINVAL;
  }
  if ((rc = check_options(view, shs, colors_precomp, scales, rotations, cov3D_precomp))) return rc;
  const int W = view->image_width, H = view->image_height;
  if (pose_state_bytes < mgs_raster_pose_state_bytes(P, W, H)) { set_error("forward_pose: pose_state too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  WsLayout L;
  if ((rc = ws_layout(P, W, H, R_cap, &L))) return rc;
  if (workspace_bytes < L.total) { set_error("workspace too small: %zu < %zu", workspace_bytes, L.total); return MGS_ENOSPC; }
  const ViewConst v = make_view_const(view);
  float* ea = reinterpret_cast<float*>(pose_state);
  float* basis = reinterpret_cast<float*>(reinterpret_cast<char*>(pose_state) + pose_deferred_bytes(W, H));
  // a fresh workspace: the folded scan wants a zero header (ticket) and zero tile counters
  MGS_CUDA_TRY(cudaMemsetAsync(workspace, 0, sizeof(WsHeader), s));
  MGS_CUDA_TRY(cudaMemsetAsync(ws_ptr<void>(workspace, L.tile_ctr), 0, (size_t)v.gx * v.gy * TILE_CTR_STRIDE * sizeof(uint32_t), s));
  const FoldScan fs{R_cap, /*heavy_first=*/true, /*in_kernel=*/false};
  if ((rc = launch_preprocess(v, P, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                              workspace, L, s, basis, 32, nullptr, &fs)))
    return rc;
  if (status_out) MGS_CUDA_TRY(cudaMemcpyAsync(status_out, workspace, 6 * sizeof(uint32_t), cudaMemcpyDefault, s));
  if (ready_event) MGS_CUDA_TRY(cudaEventRecord((cudaEvent_t)ready_event, s));
  MGS_CUDA_TRY(cudaMemsetAsync(n_touched, 0, (size_t)P * sizeof(int32_t), s));
  if ((rc = launch_binning(v, P, colors_precomp, radii, workspace, L, R_cap, s, true, /*scan_done=*/true))) return rc;
  FusedLoss none{};
  return launch_blend_track(v, workspace, L, basis, out_color, out_depth, out_opacity, n_touched, none, true, nullptr, s,
                            nullptr, 0.f, ea, nullptr, nullptr, /*deferred=*/true);
}

int mgs_raster_backward_pose_deferred(int32_t W, int32_t H, const void* pose_state, const float* dL_dout_color,
                                      const float* dL_dout_depth, float* tau_image, float* dL_dtau_sum, void* stream) {
  if (W <= 0 || H <= 0 || !pose_state || !dL_dout_color || (!tau_image && !dL_dtau_sum)) { set_error("backward_pose_deferred: bad arguments"); return MGS_EINVAL; }
  return launch_pose_contract(reinterpret_cast<const float*>(pose_state), W, H, dL_dout_color, dL_dout_depth, tau_image, dL_dtau_sum,
                              (cudaStream_t)stream);
}

// test hook: byte offsets of the workspace regions, in WsLayout order (tests read keys/ranges through it)
int mgs_debug_layout(int32_t P, int32_t W, int32_t H, int64_t R_cap, size_t* out /* >= 17 */) {
  WsLayout L;
  int rc = ws_layout(P, W, H, R_cap, &L);
  if (rc) return rc;
  const size_t v[17] = {L.header, L.depths, L.means2D, L.conic_opacity, L.rgb4, L.tiles_touched, L.tile_ctr,
                        L.final_T, L.n_contrib, L.ranges, L.records, L.tile_ctr, L.list, L.total, 0, 0, 0};
  for (int i = 0; i < 17; i++) out[i] = v[i];
  return MGS_OK;
}

size_t mgs_raster_backward_scratch_bytes(int32_t P) { return backward_scratch_bytes(P, 0); }

static int raster_backward_impl(const mgs_view* view, int32_t P, uint32_t grad_mask, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* scales, const float* rotations,
                        const float* cov3D_precomp, const int32_t* radii, const void* workspace,
                        size_t workspace_bytes, void* scratch, size_t scratch_bytes, const float* dL_dout_color,
                        const float* dL_dout_depth, float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors_or_sh,
                        float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                        float* dL_dtau_sum, float* dL_dtau_per_gaussian, const int32_t* pixel_mult, void* stream) {
  int rc = check_view(view, true);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (P < 0) { set_error("P < 0"); return MGS_EINVAL; }
  if (P == 0) {
    if (dL_dtau_sum) MGS_CUDA_TRY(cudaMemsetAsync(dL_dtau_sum, 0, 6 * sizeof(float), s));
    return MGS_OK;
  }
  if (!means3D || !radii || !workspace || !dL_dout_color || !dL_dout_depth) { set_error("null required input"); return MGS_EINVAL; }
  if ((rc = check_options(view, shs, colors_precomp, scales, rotations, cov3D_precomp))) return rc;
  if ((grad_mask & MGS_GRAD_MEANS3D) && !dL_dmeans3D) { set_error("dL_dmeans3D requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_MEANS2D) && !dL_dmeans2D) { set_error("dL_dmeans2D requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_COLOR) && !dL_dcolors_or_sh) { set_error("colour gradient requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_OPACITY) && !dL_dopacity) { set_error("dL_dopacity requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_COV) && scales && (!dL_dscales || !dL_drotations)) { set_error("dL_dscales/dL_drotations requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_COV) && cov3D_precomp && !dL_dcov3D) { set_error("dL_dcov3D requested but null"); return MGS_EINVAL; }
  if ((grad_mask & MGS_GRAD_TAU) && !dL_dtau_sum && !dL_dtau_per_gaussian) { set_error("tau gradient requested but null"); return MGS_EINVAL; }
  BackwardArgs a;
  memset(&a, 0, sizeof(a));
  a.v = make_view_const(view); a.P = P; a.grad_mask = grad_mask;
  a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp; a.scales = scales; a.rotations = rotations;
  a.cov3D_precomp = cov3D_precomp; a.radii = radii; a.ws = workspace;
  if ((rc = ws_layout(P, view->image_width, view->image_height, 0, &a.L))) return rc;
  if (workspace_bytes < a.L.total) { set_error("workspace too small"); return MGS_ENOSPC; }
  a.scratch = scratch; a.scratch_bytes = scratch_bytes;
  a.dL_dcolor_img = dL_dout_color; a.dL_ddepth_img = dL_dout_depth;
  a.dL_dmeans3D = dL_dmeans3D; a.dL_dmeans2D = dL_dmeans2D; a.dL_dcolors_or_sh = dL_dcolors_or_sh;
  a.dL_dopacity = dL_dopacity; a.dL_dscales = dL_dscales; a.dL_drotations = dL_drotations; a.dL_dcov3D = dL_dcov3D;
  a.dL_dtau_sum = (grad_mask & MGS_GRAD_TAU) ? dL_dtau_sum : nullptr;
  a.dL_dtau_pg = (grad_mask & MGS_GRAD_TAU) ? dL_dtau_per_gaussian : nullptr;
  a.pixel_mult = pixel_mult;
  return launch_backward_full(a, s);
}

int mgs_raster_backward(const mgs_view* view, int32_t P, uint32_t grad_mask, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* scales, const float* rotations,
                        const float* cov3D_precomp, const int32_t* radii, const void* workspace,
                        size_t workspace_bytes, void* scratch, size_t scratch_bytes, const float* dL_dout_color,
                        const float* dL_dout_depth, float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dcolors_or_sh,
                        float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                        float* dL_dtau_sum, float* dL_dtau_per_gaussian, void* stream) {
  return raster_backward_impl(view, P, grad_mask, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                              workspace, workspace_bytes, scratch, scratch_bytes, dL_dout_color, dL_dout_depth, dL_dmeans3D,
                              dL_dmeans2D, dL_dcolors_or_sh, dL_dopacity, dL_dscales, dL_drotations, dL_dcov3D, dL_dtau_sum,
                              dL_dtau_per_gaussian, nullptr, stream);
}

// ---- pixel-sample backward (renderPixelSampleCUDA, backward.cu:592-797; reached through select_pixels = true of
//      _C.rasterize_gaussians_backward, rasterize_points.cu:174-189).  The reference runs one 128-thread block per listed
//      pixel in which thread 0 alone walks the pixel's list.  Here the list becomes a multiplicity plane (one scatter
//      kernel) and the regular tile-parallel backward runs with it: same arithmetic per (pixel, Gaussian) pair.
size_t mgs_raster_backward_pixels_scratch_bytes(int32_t P, int32_t W, int32_t H) {
  if (P < 0 || W <= 0 || H <= 0) return 0;
  return ((mgs_raster_backward_scratch_bytes(P) + 255) & ~(size_t)255) + (size_t)W * H * sizeof(int32_t);
}

int mgs_raster_backward_pixels(const mgs_view* view, int32_t P, uint32_t grad_mask, const float* means3D, const float* shs,
                               const float* colors_precomp, const float* scales, const float* rotations,
                               const float* cov3D_precomp, const int32_t* radii, const void* workspace,
                               size_t workspace_bytes, void* scratch, size_t scratch_bytes,
                               const int32_t* selected_pixel_indices, int32_t n_selected, const float* dL_dout_color,
                               const float* dL_dout_depth, float* dL_dmeans3D, float* dL_dmeans2D,
                               float* dL_dcolors_or_sh, float* dL_dopacity, float* dL_dscales, float* dL_drotations,
                               float* dL_dcov3D, float* dL_dtau_sum, float* dL_dtau_per_gaussian, void* stream) {
  int rc = check_view(view, true);
  if (rc) return rc;
  if (n_selected < 0 || (n_selected > 0 && !selected_pixel_indices)) { set_error("backward_pixels: bad pixel list"); return MGS_EINVAL; }
  if (P > 0 && !scratch) { set_error("backward_pixels: null scratch"); return MGS_EINVAL; }
  const int W = view->image_width, H = view->image_height;
  if (P > 0 && scratch_bytes < mgs_raster_backward_pixels_scratch_bytes(P, W, H)) { set_error("backward_pixels: scratch too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  int32_t* mult = nullptr;
  size_t inner = 0;
  if (P > 0) {
    inner = (mgs_raster_backward_scratch_bytes(P) + 255) & ~(size_t)255;
    mult = reinterpret_cast<int32_t*>(static_cast<unsigned char*>(scratch) + inner);
    if ((rc = launch_pixel_multiplicity(selected_pixel_indices, n_selected, W * H, mult, s))) return rc;
  }
  return raster_backward_impl(view, P, grad_mask, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                              workspace, workspace_bytes, scratch, inner, dL_dout_color, dL_dout_depth, dL_dmeans3D,
                              dL_dmeans2D, dL_dcolors_or_sh, dL_dopacity, dL_dscales, dL_drotations, dL_dcov3D, dL_dtau_sum,
                              dL_dtau_per_gaussian, mult, stream);
}


size_t mgs_raster_backward_pose_scratch_bytes(int32_t P, int32_t W, int32_t H) { return backward_pose_scratch_bytes(P, W, H); }

int mgs_raster_backward_pose(const mgs_view* view, int32_t P, const float* means3D, const float* shs,
                             const float* colors_precomp, const float* scales, const float* rotations,
                             const float* cov3D_precomp, const int32_t* radii, const void* workspace,
                             size_t workspace_bytes, void* scratch, size_t scratch_bytes, const float* dL_dout_color,
                             const float* dL_dout_depth, float* tau_image, float* dL_dtau_sum, void* stream) {
  int rc = check_view(view, true);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (P < 0) { set_error("P < 0"); return MGS_EINVAL; }
  if (!tau_image && !dL_dtau_sum) { set_error("backward_pose: no output requested"); return MGS_EINVAL; }
  const size_t npix = (size_t)view->image_width * view->image_height;
  if (P == 0) {
    if (dL_dtau_sum) MGS_CUDA_TRY(cudaMemsetAsync(dL_dtau_sum, 0, 6 * sizeof(float), s));
    if (tau_image) MGS_CUDA_TRY(cudaMemsetAsync(tau_image, 0, 6 * npix * sizeof(float), s));
    return MGS_OK;
  }
  if (!means3D || !radii || !workspace || !dL_dout_color) { set_error("null required input"); return MGS_EINVAL; }   // dL_dout_depth may be null (= zeros)
  if ((rc = check_options(view, shs, colors_precomp, scales, rotations, cov3D_precomp))) return rc;
  BackwardArgs a;
  memset(&a, 0, sizeof(a));
  a.v = make_view_const(view); a.P = P; a.grad_mask = MGS_GRAD_TAU;
  a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp; a.scales = scales; a.rotations = rotations;
  a.cov3D_precomp = cov3D_precomp; a.radii = radii; a.ws = workspace;
  if ((rc = ws_layout(P, view->image_width, view->image_height, 0, &a.L))) return rc;
  if (workspace_bytes < a.L.total) { set_error("workspace too small"); return MGS_ENOSPC; }
  a.scratch = scratch; a.scratch_bytes = scratch_bytes;
  a.dL_dcolor_img = dL_dout_color; a.dL_ddepth_img = dL_dout_depth;
  a.dL_dtau_sum = dL_dtau_sum; a.sketch_dtau = tau_image;
  return launch_backward_pose(a, s);
}

int mgs_sketch_reduce(const float* tau_image, int32_t H, int32_t W, int32_t n_stacks, int32_t sketch_dim,
                      const int32_t* sketch_indices, float* sketch_dtau, void* stream) {
  if (!tau_image || !sketch_indices || !sketch_dtau || H <= 0 || W <= 0 || n_stacks <= 0 || sketch_dim <= 0) {
    set_error("sketch_reduce: bad arguments"); return MGS_EINVAL;
  }
  if (sketch_dim > 64) { set_error("sketch_dim must be <= 64 (__init__.py:249)"); return MGS_EINVAL; }
  return launch_sketch_reduce(tau_image, H, W, n_stacks, sketch_dim, sketch_indices, sketch_dtau, (cudaStream_t)stream, 6);
}

}  // extern "C"
