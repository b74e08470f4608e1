// mgs_track_iteration — one whole pose-tracking iteration (the inner loop of FrontEnd.tracking,
// MonoGS utils/slam_frontend.py:455-630, first-order phase) as a fixed, host-sync-free launch sequence:
//
//   preprocess (+ pose Jacobian of visible Gaussians, same thread)           1 kernel
//   depth sort, scan, header, emit keys, tile sort, ranges + records         3 kernels + CUB
//   SH degree 0: single-pass blend = forward + loss + pose tangent (blend_track.cu)   1 kernel
//   SH degree > 0: forward blend with the loss fused into its epilogue        1 kernel
//       (per-pixel residual, Huber / L1 term, un-normalised dL/dimage, per-tile partial sums)
//     + pose-only backward blend (per-pixel tau accumulation)                 1 kernel
//   update: loss + norm from the partials, tau/dab scaling, Adam (device step counter), T <- Exp(tau) T,
//           view / projection matrices                                        1 kernel
//
// The Huber/L2 objective needs ||h|| before its gradient; but the pose tangent is LINEAR in dL/dimage, so the
// backward runs on the un-normalised gradient and the update kernel applies 1/||h|| to tau and to dL/d(a,b).
// Compared with calling the individual entry points this drops 8 launches and one full read + write of the
// rendered images per iteration.
#include "launchers.h"
#include "loss_pixel.cuh"
#include "pose_jac.cuh"
#include "pose_update.cuh"
#include <cstring>

namespace mgs {

struct TrackUpdateK {
  int mode, use_depth, npix, tiles;
  float alpha, scale;
  const double* partials; const float* tau_raw;
  float *loss, *tau, *dab;
  float *params, *exp_avg, *exp_avg_sq; int* step; float lr[4]; float beta1, beta2, eps;
  float *T, *view_out, *proj_out; const float* proj_raw; int* converged; float thr;
};

__global__ void __launch_bounds__(256) track_update_kernel(const TrackUpdateK k) {
  __shared__ double sh[256][4];
  double acc[4] = {0, 0, 0, 0};
  for (int i = threadIdx.x; i < k.tiles; i += 256)
#pragma unroll
    for (int q = 0; q < 4; q++) acc[q] += k.partials[(size_t)i * 4 + q];
#pragma unroll
  for (int q = 0; q < 4; q++) sh[threadIdx.x][q] = acc[q];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o)
#pragma unroll
      for (int q = 0; q < 4; q++) sh[threadIdx.x][q] += sh[threadIdx.x + o][q];
    __syncthreads();
  }
  if (threadIdx.x != 0) return;
  float gscale = 1.f;
  if (k.mode == MGS_LOSS_TRACK_HUBER) {
    const float nrm = (float)sqrt(sh[0][0]);
    k.loss[0] = k.scale * nrm;
    gscale = nrm > 0.f ? k.scale / nrm : 0.f;
  } else {
    const double rgb = sh[0][0] / (3.0 * k.npix), dep = sh[0][3] / (double)k.npix;
    k.loss[0] = (float)(k.scale * (k.use_depth ? k.alpha * rgb + (1.0 - k.alpha) * dep : rgb));
  }
  float g[8];
#pragma unroll
  for (int i = 0; i < 6; i++) { g[i] = k.tau_raw[i] * gscale; k.tau[i] = g[i]; }
  g[6] = (float)sh[0][1] * gscale; g[7] = (float)sh[0][2] * gscale;
  k.dab[0] = g[6]; k.dab[1] = g[7];
  // Adam (torch.optim.Adam, eps outside the bias-corrected sqrt) on [rho(3), theta(3), a, b]
  const double t = (double)(k.step[0] + 1);
  k.step[0] += 1;
  const float bc1 = (float)(1.0 - pow((double)k.beta1, t)), bc2s = (float)sqrt(1.0 - pow((double)k.beta2, t));
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const float lr = k.lr[i < 3 ? 0 : (i < 6 ? 1 : i - 4)];
    float m = k.exp_avg[i], v = k.exp_avg_sq[i];
    m = m + (g[i] - m) * (1.f - k.beta1);
    v = v * k.beta2 + (1.f - k.beta2) * g[i] * g[i];
    const float denom = sqrtf(v) / bc2s + k.eps;
    k.params[i] = k.params[i] - (lr / bc1) * (m / denom);
    k.exp_avg[i] = m; k.exp_avg_sq[i] = v;
  }
  pose_update_dev(k.T, k.params, k.params + 3, k.proj_raw, k.view_out, k.proj_out, k.converged, k.thr);
}

}  // namespace mgs

using namespace mgs;

extern "C" {

size_t mgs_track_scratch_bytes(int32_t P, int32_t W, int32_t H) {
  const size_t Pp = P > 0 ? P : 1, npix = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  return up(Pp * PoseJac<9>::kFloats * 4) + up(3 * npix * 4) + up(npix * 4) + up(tiles * 4 * 8) + 256;
}

int mgs_track_iteration(const mgs_track_args* t, void* stream) {
  if (!t) { set_error("track: args is null"); return MGS_EINVAL; }
  const mgs_view* view = &t->view;
  if (view->image_height <= 0 || view->image_width <= 0 || !view->bg || !view->viewmatrix || !view->projmatrix ||
      !view->projmatrix_raw) { set_error("track: incomplete view"); return MGS_EINVAL; }
  const int P = t->P, W = view->image_width, H = view->image_height;
  if (P <= 0 || !t->means3D || !t->opacities || !t->shs || !t->scales || !t->rotations) { set_error("track: scene pointers"); return MGS_EINVAL; }
  if (!t->gt_image || !t->T_w2c || !t->params || !t->exp_avg || !t->exp_avg_sq || !t->step_counter || !t->out_color ||
      !t->out_depth || !t->out_opacity || !t->radii || !t->n_touched || !t->loss || !t->tau || !t->dab || !t->workspace ||
      !t->scratch) { set_error("track: null buffer"); return MGS_EINVAL; }
  if (t->loss_mode != MGS_LOSS_TRACK_L1 && t->loss_mode != MGS_LOSS_TRACK_HUBER) { set_error("track: loss_mode must be a tracking mode"); return MGS_EINVAL; }
  const int use_depth = t->use_depth && t->loss_mode == MGS_LOSS_TRACK_L1;
  if (use_depth && !t->gt_depth) { set_error("track: use_depth needs gt_depth"); return MGS_EINVAL; }
  if (t->R_cap <= 0 || t->R_cap > 0x7fffffffLL) { set_error("track: bad R_cap"); return MGS_EINVAL; }
  if (t->scratch_bytes < mgs_track_scratch_bytes(P, W, H)) { set_error("track: scratch too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  WsLayout L;
  int rc = ws_layout(P, W, H, t->R_cap, &L);
  if (rc) return rc;
  if (t->workspace_bytes < L.total) { set_error("track: workspace too small: %zu < %zu", t->workspace_bytes, L.total); return MGS_ENOSPC; }
  const ViewConst v = make_view_const(view);
  const size_t npix = (size_t)W * H;
  const int tiles = v.gx * v.gy;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  char* sp = reinterpret_cast<char*>(t->scratch);
  float* jac = reinterpret_cast<float*>(sp); sp += up((size_t)P * PoseJac<9>::kFloats * 4);
  float* g_img = reinterpret_cast<float*>(sp); sp += up(3 * npix * 4);
  float* g_depth = reinterpret_cast<float*>(sp); sp += up(npix * 4);
  double* partials = reinterpret_cast<double*>(sp); sp += up((size_t)tiles * 4 * 8);
  float* tau_raw = reinterpret_cast<float*>(sp);
  const bool with_color = v.D > 0;
  void* ws = t->workspace;

  if ((rc = launch_preprocess(v, P, t->means3D, t->opacities, t->shs, nullptr, t->scales, t->rotations, nullptr, t->radii,
                              ws, L, s, jac, with_color ? 9 : 6)))
    return rc;
  if (!t->fast_math) MGS_CUDA_TRY(cudaMemsetAsync(t->n_touched, 0, (size_t)P * sizeof(int32_t), s));
  MGS_CUDA_TRY(cudaMemsetAsync(tau_raw, 0, 6 * sizeof(float), s));
  if ((rc = launch_binning(v, P, nullptr, t->radii, ws, L, t->R_cap, s))) return rc;

  LossK lk;
  lk.mode = t->loss_mode; lk.npix = (int)npix; lk.use_depth = use_depth;
  lk.thr = t->rgb_boundary_threshold; lk.alpha = t->alpha; lk.delta = t->huber_delta; lk.scale = 1.f; lk.eps = 1e-8f;
  lk.image = nullptr; lk.depth = nullptr; lk.opacity = nullptr; lk.gt = t->gt_image; lk.gt_depth = t->gt_depth;
  lk.mask = t->grad_mask; lk.a = t->params + 6; lk.b = t->params + 7;
  FusedLoss fl;
  fl.k = &lk;
  fl.grad_scale = (t->loss_mode == MGS_LOSS_TRACK_HUBER) ? 1.f : (use_depth ? t->alpha : 1.f) / (3.f * (float)npix);
  fl.g_img = g_img; fl.g_depth = use_depth ? g_depth : nullptr; fl.partials = partials; fl.fast = t->fast_math != 0;
  if (!with_color) {
    // SH degree 0: colours do not depend on the pose -> forward, loss and pose tangent in one pass over the tile lists
    if ((rc = launch_blend_track(v, ws, L, jac, t->out_color, t->out_depth, t->out_opacity, t->n_touched, fl, use_depth != 0,
                                 tau_raw, s)))
      return rc;
  } else {
    if ((rc = launch_blend_forward(v, ws, L, t->out_color, t->out_depth, t->out_opacity, t->n_touched, s, &fl))) return rc;
    BackwardArgs a;
    memset(&a, 0, sizeof(a));
    a.v = v; a.P = P; a.grad_mask = MGS_GRAD_TAU; a.ws = ws; a.L = L;
    a.dL_dcolor_img = g_img; a.dL_ddepth_img = use_depth ? g_depth : nullptr; a.dL_dtau_sum = tau_raw;
    if ((rc = launch_blend_backward_pose(a, jac, with_color, s, t->fast_math != 0))) return rc;
  }

  TrackUpdateK k;
  k.mode = t->loss_mode; k.use_depth = use_depth; k.npix = (int)npix; k.tiles = tiles; k.alpha = t->alpha; k.scale = 1.f;
  k.partials = partials; k.tau_raw = tau_raw; k.loss = t->loss; k.tau = t->tau; k.dab = t->dab;
  k.params = t->params; k.exp_avg = t->exp_avg; k.exp_avg_sq = t->exp_avg_sq; k.step = t->step_counter;
  for (int i = 0; i < 4; i++) k.lr[i] = t->lr[i];
  k.beta1 = t->beta1; k.beta2 = t->beta2; k.eps = t->adam_eps;
  k.T = t->T_w2c; k.view_out = const_cast<float*>(view->viewmatrix); k.proj_out = const_cast<float*>(view->projmatrix);
  k.proj_raw = view->projmatrix_raw; k.converged = t->converged; k.thr = t->converged_threshold;
  MGS_PROF("track_update", s, track_update_kernel<<<1, 256, 0, s>>>(k));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
