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
#include "track_update.cuh"
#include <cstring>
#include <cstdlib>

namespace mgs {

int launch_sketch_reduce(const float* rows, int H, int W, int n_stacks, int d, const int32_t* sketch_indices,
                         float* out, cudaStream_t s, int nv);

__global__ void __launch_bounds__(256) track_update_kernel(const TrackUpdateK k) { pdl_enter<PDL_UPD>(); track_update_body(k); }

// ---- second-order (sketched Gauss-Newton / Levenberg-Marquardt) iteration, utils/slam_frontend.py:455-720 with
//      in_second_order == True.  The reference applies the step of iteration k at the START of iteration k+1
//      (new_viewpoint_params.assign + update_pose, :474-480), renders, adapts lambda on the new scalar loss (:541-548),
//      builds SJ / Sf through `stack_dim` sketch-Jacobian backward passes, solves
//      lstsq([SJ; sqrt(lambda) I], -[Sf; 0]) (:668-679) and leaves x pending.  Same order here, three small kernels
//      around the rasterizer: begin (apply the pending step), [preprocess .. blend_track<SKETCH> .. sketch_reduce<9>],
//      solve (scalar loss, lambda, 8x8 normal equations in double, Cholesky, pending x, best-pose record). ----
struct Track2K {
  int npix, tiles, rows;
  const double* partials; const float* sj;   // sj: f32[rows][9] = 8 Jacobian columns (rho, theta, a, b) | Sf
  float *params, *T, *view_out, *proj_out; const float* proj_raw;
  float* st;
  float min_lambda, max_lambda, inc, dec, conv_thr;
  float* loss_out;   // f32[1]: the scalar ||r||_1 of this iteration
};

__global__ void __launch_bounds__(32) track2_begin_kernel(const Track2K k) {
  __shared__ float s_T[16], s_p[8], s_view[16], s_proj[16], s_pr[16];
  pdl_enter<PDL_UPD>();
  const int tid = threadIdx.x;
  if (k.st[LM_PENDING] == 0.f) return;     // uniform
  if (tid < 16) { s_T[tid] = k.T[tid]; s_pr[tid] = k.proj_raw[tid]; }
  if (tid < 8) s_p[tid] = k.params[tid] + k.st[LM_X + tid];    // TempCamera.step (slam_frontend.py:49-53)
  __syncwarp();
  if (tid == 0) pose_update_dev(s_T, s_p, s_p + 3, s_pr, s_view, s_proj, nullptr, 0.f);   // update_pose: T <- Exp(tau) T, deltas = 0
  __syncwarp();
  if (tid < 16) { k.T[tid] = s_T[tid]; k.view_out[tid] = s_view[tid]; k.proj_out[tid] = s_proj[tid]; }
  if (tid < 8) k.params[tid] = s_p[tid];
  if (tid == 0) k.st[LM_PENDING] = 0.f;
}

__global__ void __launch_bounds__(256) track2_solve_kernel(const Track2K k) {
  constexpr int NA = 36 + 8;                  // upper triangle of J^T J, then J^T f
  __shared__ double red[8][NA + 1];
  __shared__ double tot[NA + 1];
  pdl_enter<PDL_UPD>();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double acc[NA + 1];
#pragma unroll
  for (int q = 0; q <= NA; q++) acc[q] = 0.0;
  for (int r = tid; r < k.rows; r += 256) {
    float j[9];
#pragma unroll
    for (int c = 0; c < 9; c++) j[c] = k.sj[(size_t)r * 9 + c];
    int q = 0;
#pragma unroll
    for (int a = 0; a < 8; a++)
#pragma unroll
      for (int b = a; b < 8; b++) acc[q++] += (double)j[a] * (double)j[b];
#pragma unroll
    for (int a = 0; a < 8; a++) acc[36 + a] += (double)j[a] * (double)j[8];
  }
  for (int i = tid; i < k.tiles; i += 256) acc[NA] += k.partials[(size_t)i * 4];   // sum |r| of the un-Hubered residual
#pragma unroll
  for (int q = 0; q <= NA; q++) {
    double x = acc[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) red[warp][q] = x;
  }
  __syncthreads();
  if (tid <= NA) {
    double x = 0.0;
#pragma unroll
    for (int w = 0; w < 8; w++) x += red[w][tid];
    tot[tid] = x;
  }
  __syncthreads();
  if (tid != 0) return;
  float* st = k.st;
  const float scalar = (float)tot[NA];
  // lambda: a step that lowered the loss relaxes the damping, one that raised it tightens it (:541-548); the step
  // itself is always kept (the reference forces is_new_step = True, :551)
  float lambda = st[LM_LAMBDA];
  if (st[LM_HAVE_PREV] != 0.f) {
    if (scalar < st[LM_PREV_SCALAR]) lambda = fmaxf(lambda / k.dec, k.min_lambda);
    else lambda = fminf(lambda * k.inc, k.max_lambda);
  }
  st[LM_LAMBDA] = lambda; st[LM_PREV_SCALAR] = scalar; st[LM_HAVE_PREV] = 1.f;
  lm_track_best(st, scalar, k.T, k.params[6], k.params[7]);
  // (J^T J + lambda I) x = -J^T f  ==  argmin ||[J; sqrt(lambda) I] x + [f; 0]||  (torch.linalg.lstsq, :679), in double
  double A[8][8], g[8], x[8];
  int q = 0;
  for (int a = 0; a < 8; a++) for (int b = a; b < 8; b++) { A[a][b] = A[b][a] = tot[q++]; }
  for (int a = 0; a < 8; a++) { A[a][a] += (double)lambda; g[a] = -tot[36 + a]; }
  bool ok = true;
  for (int a = 0; a < 8 && ok; a++) {          // Cholesky A = L L^T, L stored in the lower triangle
    for (int b = 0; b <= a; b++) {
      double sum = A[a][b];
      for (int c = 0; c < b; c++) sum -= A[a][c] * A[b][c];
      if (a == b) { if (!(sum > 0.0)) { ok = false; break; } A[a][a] = sqrt(sum); }
      else A[a][b] = sum / A[b][b];
    }
  }
  if (ok) {
    for (int a = 0; a < 8; a++) { double sum = g[a]; for (int c = 0; c < a; c++) sum -= A[a][c] * x[c]; x[a] = sum / A[a][a]; }
    for (int a = 7; a >= 0; a--) { double sum = x[a]; for (int c = a + 1; c < 8; c++) sum -= A[c][a] * x[c]; x[a] = sum / A[a][a]; }
  } else for (int a = 0; a < 8; a++) x[a] = 0.0;
  double n2 = 0.0;
  for (int a = 0; a < 8; a++) { st[LM_X + a] = (float)x[a]; n2 += x[a] * x[a]; }
  st[LM_PENDING] = 1.f;
  st[LM_STEP_NORM] = (float)sqrt(n2);
  st[LM_CONVERGED] = sqrt(n2) < (double)k.conv_thr ? 1.f : 0.f;   // second_order_converged (:688)
  st[LM_ITERS] += 1.f;
  if (k.loss_out) k.loss_out[0] = scalar;
}

}  // namespace mgs

using namespace mgs;

extern "C" {

size_t mgs_track_scratch_bytes(int32_t P, int32_t W, int32_t H) {
  const size_t Pp = P > 0 ? P : 1, npix = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  return up(Pp * PoseJac<9>::kFloats * 4) + up(3 * npix * 4) + up(npix * 4) + up(2 * tiles * 4 * 8) + 256 +
         up(9 * npix * 4);   // + the per-pixel Jacobian rows of the second-order iteration
}

// Packed-row staging of the single-pass blend (blend_track_kernel<ROWS>, pose_jac.cuh): preprocess writes one 176-byte row per
// Gaussian, the sort kernel writes sorted ids only, the blend gathers rows by id.  MGS_TRACK_ROWS=0 selects the per-instance
// 48-byte records + separate basis gather (the comparison is in profiles/r02_notes.txt).
static bool track_update_fold() {
  static const bool on = [] { const char* e = getenv("MGS_UPDATE_FOLD"); return e && e[0] == '1'; }();
  return on;
}
static bool track_rows() {
  static const bool on = [] { const char* e = getenv("MGS_TRACK_ROWS"); return !(e && e[0] == '0'); }();
  return on;
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
  PdlScope pdl_scope;     // the iteration's kernels are chained by programmatic dependent launch (common.cuh)
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
  double* partials = reinterpret_cast<double*>(sp); sp += up((size_t)2 * tiles * 4 * 8);
  float* tau_raw = reinterpret_cast<float*>(sp);
  const bool with_color = v.D > 0;
  void* ws = t->workspace;
  if (t->status_out) MGS_CUDA_TRY(memcpy_async(t->status_out, ws, 6 * sizeof(uint32_t), cudaMemcpyDefault, s));   // of the PREVIOUS iteration

  // The workspace and the scratch of a tracker are zero-filled ONCE by their owner: tile counters, the scan ticket, the
  // tangent accumulator and the update ticket are all left zero by the iteration that used them (the sort kernel clears
  // its tile's counter, the update kernel re-zeroes tau_raw), so an iteration is 7 kernels and no memset.
  const bool rows = !with_color && track_rows() && !track_update_fold();   // (the folded-update kernels keep the record staging)
  // tile_segments: fixed 1024-instance segment per tile, scattered by the preprocess kernel and sorted in place — the
  // iteration is preprocess, tile_sort_segments, blend_track, track_update (four kernels instead of seven)
  const bool segments = t->tile_segments != 0;
  if (segments && !rows) { set_error("track: tile_segments needs sh_degree 0 and the packed-row staging (MGS_TRACK_ROWS, MGS_UPDATE_FOLD unset)"); return MGS_EINVAL; }
  if (segments && (int64_t)tiles * TILE_SEGMENT > t->R_cap) { set_error("track: tile_segments needs R_cap >= tiles x %d", TILE_SEGMENT); return MGS_ENOSPC; }
  const FoldScan fs{t->R_cap, /*heavy_first=*/!with_color, /*in_kernel=*/false, segments ? TILE_SEGMENT : 0};
  if ((rc = launch_preprocess(v, P, t->means3D, t->opacities, t->shs, nullptr, t->scales, t->rotations, nullptr, t->radii,
                              ws, L, s, jac, with_color ? 9 : (rows ? TRACK_ROW_FLOATS : 32), nullptr, &fs)))
    return rc;
  if (!t->fast_math) MGS_CUDA_TRY(memset_async(t->n_touched, 0, (size_t)P * sizeof(int32_t), s));
  if (with_color) MGS_CUDA_TRY(memset_async(tau_raw, 0, 6 * sizeof(float), s));
  if (segments) { if ((rc = launch_sort_segments(v, ws, L, t->R_cap, s, /*ids_only=*/true))) return rc; }
  else if ((rc = launch_binning(v, P, nullptr, t->radii, ws, L, t->R_cap, s, fs.heavy_first, /*scan_done=*/true, /*clear_counts=*/true,
                                /*ids_only=*/rows))) return rc;

  LossK lk;
  lk.mode = t->loss_mode; lk.npix = (int)npix; lk.use_depth = use_depth;
  lk.thr = t->rgb_boundary_threshold; lk.alpha = t->alpha; lk.delta = t->huber_delta; lk.scale = 1.f; lk.eps = 1e-8f;
  lk.image = nullptr; lk.depth = nullptr; lk.opacity = nullptr; lk.gt = t->gt_image; lk.gt_depth = t->gt_depth;
  lk.mask = t->grad_mask; lk.a = t->params + 6; lk.b = t->params + 7;
  FusedLoss fl;
  fl.k = &lk;
  fl.grad_scale = (t->loss_mode == MGS_LOSS_TRACK_HUBER) ? 1.f : (use_depth ? t->alpha : 1.f) / (3.f * (float)npix);
  fl.g_img = g_img; fl.g_depth = use_depth ? g_depth : nullptr; fl.partials = partials; fl.fast = t->fast_math != 0;
  TrackUpdateK k;
  k.mode = t->loss_mode; k.use_depth = use_depth; k.npix = (int)npix; k.tiles = with_color ? tiles : tiles * (8 / blend_track_warps()); k.alpha = t->alpha; k.scale = 1.f;
  k.partials = partials; k.tau_raw = tau_raw; k.loss = t->loss; k.tau = t->tau; k.dab = t->dab;
  k.params = t->params; k.exp_avg = t->exp_avg; k.exp_avg_sq = t->exp_avg_sq; k.step = t->step_counter;
  for (int i = 0; i < 4; i++) k.lr[i] = t->lr[i];
  k.beta1 = t->beta1; k.beta2 = t->beta2; k.eps = t->adam_eps;
  k.T = t->T_w2c; k.view_out = const_cast<float*>(view->viewmatrix); k.proj_out = const_cast<float*>(view->projmatrix);
  k.proj_raw = view->projmatrix_raw; k.converged = t->converged; k.thr = t->converged_threshold;
  k.lm_state = t->lm_state; k.tau_clear = nullptr;
  if (!with_color) {
    // SH degree 0: colours do not depend on the pose -> forward, loss and pose tangent in one pass over the tile lists,
    // and the iteration's serial tail (loss norm, Adam, pose update) in that kernel's last CTA
    uint32_t* ticket = reinterpret_cast<uint32_t*>(tau_raw + 8);
    k.tau_clear = tau_raw;      // the update leaves the accumulator zero for the next iteration: no memset
    // The serial tail (loss norm, Adam, pose update) runs as its own one-CTA launch by default: measured on the B200 the
    // blend kernel is 24 us shorter without the per-CTA fence + ticket and the iteration 15 us shorter in total
    // (profiles/r02_notes.txt).  MGS_UPDATE_FOLD=1 selects the variant that runs it in the blend kernel's last CTA.
    const bool update_fold = track_update_fold();
    if (update_fold)
      return launch_blend_track(v, ws, L, jac, t->out_color, t->out_depth, t->out_opacity, t->n_touched, fl, use_depth != 0,
                                tau_raw, s, nullptr, 0.f, nullptr, &k, ticket);
    if ((rc = launch_blend_track(v, ws, L, jac, t->out_color, t->out_depth, t->out_opacity, t->n_touched, fl, use_depth != 0,
                                 tau_raw, s, nullptr, 0.f, nullptr, nullptr, nullptr, false, rows, segments))) return rc;
    cudaError_t e;
    MGS_PROF("track_update", s, e = launch_chain(track_update_kernel, 1, 256, 0, s, k));
    MGS_CUDA_TRY(e);
    MGS_LAUNCH_CHECK();
    return MGS_OK;
  }
  if ((rc = launch_blend_forward(v, ws, L, t->out_color, t->out_depth, t->out_opacity, t->fast_math ? nullptr : t->n_touched, s, &fl))) return rc;
  BackwardArgs a;
  memset(&a, 0, sizeof(a));
  a.v = v; a.P = P; a.grad_mask = MGS_GRAD_TAU; a.ws = ws; a.L = L;
  a.dL_dcolor_img = g_img; a.dL_ddepth_img = use_depth ? g_depth : nullptr; a.dL_dtau_sum = tau_raw;
  if ((rc = launch_blend_backward_pose(a, jac, with_color, s, t->fast_math != 0))) return rc;
  cudaError_t e;
  MGS_PROF("track_update", s, e = launch_chain(track_update_kernel, 1, 256, 0, s, k));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_track_second_order(const mgs_track2_args* a, void* stream) {
  if (!a) { set_error("track2: args is null"); return MGS_EINVAL; }
  const mgs_track_args* t = &a->t;
  const mgs_view* view = &t->view;
  if (view->image_height <= 0 || view->image_width <= 0 || !view->bg || !view->viewmatrix || !view->projmatrix ||
      !view->projmatrix_raw) { set_error("track2: incomplete view"); return MGS_EINVAL; }
  const int P = t->P, W = view->image_width, H = view->image_height;
  if (P <= 0 || !t->means3D || !t->opacities || !t->shs || !t->scales || !t->rotations) { set_error("track2: scene pointers"); return MGS_EINVAL; }
  if (!t->gt_image || !t->T_w2c || !t->params || !t->out_color || !t->out_depth || !t->out_opacity || !t->radii || !t->n_touched ||
      !t->loss || !t->workspace || !t->scratch || !t->lm_state) { set_error("track2: null buffer"); return MGS_EINVAL; }
  if (view->sh_degree != 0) { set_error("track2: the fused second-order iteration supports sh_degree 0 (the single-pass blend)"); return MGS_EINVAL; }
  if (a->n_stacks <= 0 || a->sketch_dim <= 0 || a->sketch_dim > 64 || !a->sketch_indices || !a->rand_weights || !a->rows) {
    set_error("track2: sketch arguments (sketch_dim <= 64, __init__.py:249)"); return MGS_EINVAL;
  }
  if (t->R_cap <= 0 || t->R_cap > 0x7fffffffLL) { set_error("track2: bad R_cap"); return MGS_EINVAL; }
  if (t->scratch_bytes < mgs_track_scratch_bytes(P, W, H)) { set_error("track2: scratch too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  PdlScope pdl_scope;
  WsLayout L;
  int rc = ws_layout(P, W, H, t->R_cap, &L);
  if (rc) return rc;
  if (t->workspace_bytes < L.total) { set_error("track2: workspace too small: %zu < %zu", t->workspace_bytes, L.total); return MGS_ENOSPC; }
  const ViewConst v = make_view_const(view);
  const size_t npix = (size_t)W * H;
  const int tiles = v.gx * v.gy;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  char* sp = reinterpret_cast<char*>(t->scratch);
  float* jac = reinterpret_cast<float*>(sp); sp += up((size_t)P * PoseJac<9>::kFloats * 4);
  sp += up(3 * npix * 4) + up(npix * 4);
  double* partials = reinterpret_cast<double*>(sp); sp += up((size_t)2 * tiles * 4 * 8);
  sp += 256;
  float* sk_image = reinterpret_cast<float*>(sp);
  void* ws = t->workspace;
  if (t->status_out) MGS_CUDA_TRY(memcpy_async(t->status_out, ws, 6 * sizeof(uint32_t), cudaMemcpyDefault, s));

  Track2K k;
  k.npix = (int)npix; k.tiles = tiles * (8 / blend_track_warps()); k.rows = a->n_stacks * a->sketch_dim;
  k.partials = partials; k.sj = a->rows; k.params = t->params; k.T = t->T_w2c;
  k.view_out = const_cast<float*>(view->viewmatrix); k.proj_out = const_cast<float*>(view->projmatrix);
  k.proj_raw = view->projmatrix_raw; k.st = a->lm_state ? a->lm_state : t->lm_state;
  k.min_lambda = a->min_lambda; k.max_lambda = a->max_lambda; k.inc = a->increase_factor; k.dec = a->decrease_factor;
  k.conv_thr = a->converged_threshold; k.loss_out = t->loss;
  cudaError_t e;
  // the sketch rows are accumulated by the blend kernel's epilogue (SketchRoute): cleared ahead of the chain's first kernel
  static const bool route_env = [] { const char* e = getenv("MGS_SKETCH_IN_BLEND"); return !(e && e[0] == '0'); }();
  const SketchRoute route{a->sketch_indices, a->rows, a->n_stacks, a->sketch_dim};
  if (route_env) MGS_CUDA_TRY(memset_async(a->rows, 0, (size_t)k.rows * 9 * sizeof(float), s));
  MGS_PROF("track2_begin", s, e = launch_chain(track2_begin_kernel, 1, 32, 0, s, k));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  const bool rows = track_rows();
  const bool segments = t->tile_segments != 0;
  if (segments && !rows) { set_error("track2: tile_segments needs the packed-row staging"); return MGS_EINVAL; }
  if (segments && (int64_t)tiles * TILE_SEGMENT > t->R_cap) { set_error("track2: tile_segments needs R_cap >= tiles x %d", TILE_SEGMENT); return MGS_ENOSPC; }
  const FoldScan fs{t->R_cap, /*heavy_first=*/true, /*in_kernel=*/false, segments ? TILE_SEGMENT : 0};
  if ((rc = launch_preprocess(v, P, t->means3D, t->opacities, t->shs, nullptr, t->scales, t->rotations, nullptr, t->radii,
                              ws, L, s, jac, rows ? TRACK_ROW_FLOATS : 32, nullptr, &fs)))
    return rc;
  if (!t->fast_math) MGS_CUDA_TRY(memset_async(t->n_touched, 0, (size_t)P * sizeof(int32_t), s));
  if (segments) { if ((rc = launch_sort_segments(v, ws, L, t->R_cap, s, /*ids_only=*/true))) return rc; }
  else if ((rc = launch_binning(v, P, nullptr, t->radii, ws, L, t->R_cap, s, true, /*scan_done=*/true, /*clear_counts=*/true, /*ids_only=*/rows))) return rc;
  LossK lk;
  lk.mode = MGS_LOSS_TRACK_HUBER; lk.npix = (int)npix; lk.use_depth = 0;
  lk.thr = t->rgb_boundary_threshold; lk.alpha = t->alpha; lk.delta = t->huber_delta; lk.scale = 1.f; lk.eps = 1e-8f;
  lk.image = nullptr; lk.depth = nullptr; lk.opacity = nullptr; lk.gt = t->gt_image; lk.gt_depth = nullptr;
  lk.mask = t->grad_mask; lk.a = t->params + 6; lk.b = t->params + 7;
  FusedLoss fl;
  fl.k = &lk; fl.grad_scale = 1.f; fl.g_img = nullptr; fl.g_depth = nullptr; fl.partials = partials; fl.fast = t->fast_math != 0;
  const float sk_scale = (float)((double)k.rows / (double)npix);     // 1 / (m / (stack_dim * sketch_dim)), slam_frontend.py:642
  if ((rc = launch_blend_track(v, ws, L, jac, t->out_color, t->out_depth, t->out_opacity, t->n_touched, fl, false, nullptr, s,
                               a->rand_weights, sk_scale, sk_image, nullptr, nullptr, false, rows, segments, route_env ? &route : nullptr)))
    return rc;
  if (!route_env && (rc = launch_sketch_reduce(sk_image, H, W, a->n_stacks, a->sketch_dim, a->sketch_indices, a->rows, s, 9))) return rc;
  MGS_PROF("track2_solve", s, e = launch_chain(track2_solve_kernel, 1, 256, 0, s, k));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
