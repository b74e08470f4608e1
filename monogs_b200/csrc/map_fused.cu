// Keyframe-window mapping iteration on the raw Gaussian parameters (include/monogs_b200.h: mgs_map_*).
//
// Reference flow per iteration (utils/slam_backend.py:157-333): for every view of the window, render() re-activates
// the parameters (sigmoid / exp / normalize / cat, gaussian_model.py:77-97), rasterizes, evaluates get_loss_mapping
// (utils/slam_utils.py:224-253) — then ONE autograd backward walks all of it again, and Python loops update the
// densification statistics.  Here:
//   mgs_map_begin      1 kernel : activate all parameters once, zero the gradient accumulators
//   mgs_map_view x V   preprocess -> counting-sort binning -> forward blend with the mapping loss fused into its epilogue
//                      -> full backward blend -> per-Gaussian backward that ACCUMULATES activated-space gradients and
//                      updates xyz_gradient_accum / denom / max_radii2D in the same pass.  No host synchronisation.
//   mgs_map_end        1 kernel : activation chain rule (+ the isotropic regulariser) -> raw-parameter gradients
// so the per-view cost is the rasterizer alone and the activation / regulariser / statistics work is O(P) per
// iteration instead of O(P) torch kernels per view.
#include "launchers.h"
#include "loss_pixel.cuh"
#include "pose_update.cuh"
#include <cstring>

namespace mgs {

constexpr int MAP_SLOTS = 4;

struct MapScratch {
  float *act_opacity, *act_scales, *act_rot;             // activated parameters (SH features have no activation and are
                                                         // read in place from f_dc / f_rest: split layout of sh_stage.cuh)
  float *d_opacity, *d_scales, *d_rot;                   // activated-space gradient accumulators
  void* bwd[MAP_SLOTS]; size_t bwd_bytes;                // per-Gaussian screen-space accumulator of the backward blend, per slot
  float *g_img[MAP_SLOTS], *g_depth[MAP_SLOTS]; double* partials[MAP_SLOTS];   // forwards of later views may overlap the backward of an earlier one
  size_t total;
};

static inline size_t up256(size_t x) { return (x + 255) / 256 * 256; }

static MapScratch carve_map_scratch(void* base, int P, int M, int W, int H) {
  MapScratch m;
  const size_t Pp = P > 0 ? P : 1, npix = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  char* p = reinterpret_cast<char*>(base);
  size_t off = 0;
  auto take = [&](size_t bytes) { char* q = p ? p + off : nullptr; off += up256(bytes); return q; };
  m.act_opacity = (float*)take(Pp * 4); m.act_scales = (float*)take(Pp * 12); m.act_rot = (float*)take(Pp * 16);
  m.d_opacity = (float*)take(Pp * 4); m.d_scales = (float*)take(Pp * 12); m.d_rot = (float*)take(Pp * 16);
  m.bwd_bytes = backward_scratch_bytes(P, 0);
  for (int k = 0; k < MAP_SLOTS; k++) m.bwd[k] = take(m.bwd_bytes);
  for (int k = 0; k < MAP_SLOTS; k++) {
    m.g_img[k] = (float*)take(3 * npix * 4); m.g_depth[k] = (float*)take(npix * 4);
    m.partials[k] = (double*)take(tiles * 4 * 8);
  }
  m.total = off;
  return m;
}

// sigmoid / exp / normalize as torch evaluates them (gaussian_model.py:56-63, 77-97) + zero-fill of the accumulators
__global__ void __launch_bounds__(256)
map_activate_kernel(const mgs_gaussians g, const MapScratch m, float* __restrict__ grad_xyz, float* __restrict__ grad_f_dc,
                    float* __restrict__ loss_total) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx == 0) loss_total[0] = 0.f;
  if (idx >= g.P) return;
  m.act_opacity[idx] = 1.f / (1.f + expf(-g.opacity_raw[idx]));
  float s[3];
  if (g.isotropic) s[0] = s[1] = s[2] = expf(g.scaling_raw[idx]);
  else { s[0] = expf(g.scaling_raw[3 * idx]); s[1] = expf(g.scaling_raw[3 * idx + 1]); s[2] = expf(g.scaling_raw[3 * idx + 2]); }
  m.act_scales[3 * idx] = s[0]; m.act_scales[3 * idx + 1] = s[1]; m.act_scales[3 * idx + 2] = s[2];
  const float4 q = reinterpret_cast<const float4*>(g.rotation_raw)[idx];
  const float nrm = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);   // F.normalize eps
  reinterpret_cast<float4*>(m.act_rot)[idx] = make_float4(q.x / nrm, q.y / nrm, q.z / nrm, q.w / nrm);
  if (g.sh_coeffs == 1) {   // M > 1: grad_f_dc / grad_f_rest are zeroed by memsets in mgs_map_begin
    grad_f_dc[3 * idx] = 0.f; grad_f_dc[3 * idx + 1] = 0.f; grad_f_dc[3 * idx + 2] = 0.f;
  }
  grad_xyz[3 * idx] = 0.f; grad_xyz[3 * idx + 1] = 0.f; grad_xyz[3 * idx + 2] = 0.f;
  m.d_opacity[idx] = 0.f;
  m.d_scales[3 * idx] = 0.f; m.d_scales[3 * idx + 1] = 0.f; m.d_scales[3 * idx + 2] = 0.f;
  reinterpret_cast<float4*>(m.d_rot)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
}


// per-view loss from the per-tile partial sums of the forward epilogue (get_loss_mapping_rgb / _rgbd)
__global__ void __launch_bounds__(256)
map_loss_finalize_kernel(const double* __restrict__ partials, const int tiles, const int npix, const int use_depth,
                         const float alpha, float* __restrict__ loss_view, float* __restrict__ loss_total,
                         float* __restrict__ dab) {
  __shared__ double sh[256][4];
  pdl_enter<PDL_LOSSF>();
  double acc[4] = {0, 0, 0, 0};
  for (int i = threadIdx.x; i < tiles; i += 256)
#pragma unroll
    for (int q = 0; q < 4; q++) acc[q] += partials[(size_t)i * 4 + q];
#pragma unroll
  for (int q = 0; q < 4; q++) sh[threadIdx.x][q] = acc[q];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o)
#pragma unroll
      for (int q = 0; q < 4; q++) sh[threadIdx.x][q] += sh[threadIdx.x + o][q];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double rgb = sh[0][0] / (3.0 * npix), dep = sh[0][3] / (double)npix;
    const float l = (float)(use_depth ? alpha * rgb + (1.0 - alpha) * dep : rgb);
    if (loss_view) loss_view[0] = l;
    atomicAdd(loss_total, l);   // forwards of different views may run on different streams
    if (dab) { dab[0] = (float)sh[0][1]; dab[1] = (float)sh[0][2]; }
  }
}

// activation chain rule + isotropic regulariser  weight * mean|s - mean(s)|  (slam_backend.py:244-246)
__global__ void __launch_bounds__(256)
map_end_kernel(const mgs_gaussians g, const MapScratch m, float* __restrict__ grad_f_dc, float* __restrict__ grad_f_rest,
               float* __restrict__ grad_opacity, float* __restrict__ grad_scaling, float* __restrict__ grad_rotation,
               const float reg_weight /* 0 = no regulariser on this rank */, float* __restrict__ loss_total) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float reg = 0.f;
  if (idx < g.P) {
    const float o = m.act_opacity[idx];
    grad_opacity[idx] = m.d_opacity[idx] * ((1.f - o) * o);            // sigmoid_backward: grad * (1 - y) * y
    const float s[3] = {m.act_scales[3 * idx], m.act_scales[3 * idx + 1], m.act_scales[3 * idx + 2]};
    float ds[3] = {m.d_scales[3 * idx], m.d_scales[3 * idx + 1], m.d_scales[3 * idx + 2]};
    if (reg_weight != 0.f && !g.isotropic) {
      const float mean = (s[0] + s[1] + s[2]) / 3.f;
      float sg[3], ssum = 0.f;
#pragma unroll
      for (int j = 0; j < 3; j++) { const float d = s[j] - mean; reg += fabsf(d); sg[j] = sgn(d); ssum += sg[j]; }
      const float c = reg_weight / (3.f * (float)g.P);
#pragma unroll
      for (int j = 0; j < 3; j++) ds[j] += c * (sg[j] - ssum / 3.f);
      reg *= c;
    }
    if (g.isotropic) grad_scaling[idx] = (ds[0] + ds[1] + ds[2]) * s[0];       // repeat(1,3) backward = sum; exp' = y
    else { grad_scaling[3 * idx] = ds[0] * s[0]; grad_scaling[3 * idx + 1] = ds[1] * s[1]; grad_scaling[3 * idx + 2] = ds[2] * s[2]; }
    const float4 q = reinterpret_cast<const float4*>(g.rotation_raw)[idx];
    const float nrm = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    const float4 y = reinterpret_cast<const float4*>(m.act_rot)[idx];
    const float4 d = reinterpret_cast<const float4*>(m.d_rot)[idx];
    const float yd = y.x * d.x + y.y * d.y + y.z * d.z + y.w * d.w;
    reinterpret_cast<float4*>(grad_rotation)[idx] =
        make_float4((d.x - y.x * yd) / nrm, (d.y - y.y * yd) / nrm, (d.z - y.z * yd) / nrm, (d.w - y.w * yd) / nrm);
  }
  if (reg_weight != 0.f) {   // uniform branch
    __shared__ float part[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) reg += __shfl_xor_sync(0xffffffffu, reg, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = reg;
    __syncthreads();
    if (threadIdx.x == 0) {
      float x = 0.f;
      for (int w = 0; w < 8; w++) x += part[w];
      if (x != 0.f) atomicAdd(loss_total, x);
    }
  }
}

// Window poses: keyframe_optimizers.step() + update_pose() of BackEnd.map (utils/slam_backend.py:325-332) for the views a
// rank owns.  One CTA per view; thread 0 runs the 8-parameter Adam (torch.optim.Adam, eps outside the bias-corrected
// sqrt) and T <- Exp(rho, theta) T, then the view / projection matrices go out in parallel.
constexpr int POSE_SLOTS_MAX = 16;
struct PoseSlots { mgs_pose_slot s[POSE_SLOTS_MAX]; const float* proj_raw; float beta1, beta2, eps; };

__global__ void __launch_bounds__(32) window_pose_step_kernel(const PoseSlots k) {
  const mgs_pose_slot& p = k.s[blockIdx.x];
  __shared__ float sT[16], sP[8], sView[16], sProj[16], sPr[16];
  const int tid = threadIdx.x;
  if (tid < 16) { sT[tid] = p.T_w2c[tid]; sPr[tid] = k.proj_raw[tid]; }
  if (tid < 8) sP[tid] = p.params[tid];
  __syncwarp();
  if (tid == 0 && !p.frozen) {
    const double t = (double)(p.step[0] + 1);
    p.step[0] += 1;
    const float bc1 = (float)(1.0 - pow((double)k.beta1, t)), bc2s = (float)sqrt(1.0 - pow((double)k.beta2, t));
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const float g = i < 6 ? p.tau[i] : p.dab[i - 6];
      const float lr = p.lr[i < 3 ? 0 : (i < 6 ? 1 : i - 4)];
      float m = p.exp_avg[i], v = p.exp_avg_sq[i];
      m = m + (g - m) * (1.f - k.beta1);
      v = v * k.beta2 + (1.f - k.beta2) * g * g;
      const float denom = sqrtf(v) / bc2s + k.eps;
      sP[i] = sP[i] - (lr / bc1) * (m / denom);
      p.exp_avg[i] = m; p.exp_avg_sq[i] = v;
    }
    pose_update_dev(sT, sP, sP + 3, sPr, sView, sProj, nullptr, 0.f);
  } else if (tid == 0) {
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) sView[4 * i + j] = sT[4 * j + i];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) {
      float acc = 0.f;
      for (int q = 0; q < 4; q++) acc += sT[4 * q + i] * sPr[4 * q + j];
      sProj[4 * i + j] = acc;
    }
  }
  __syncwarp();
  if (tid < 16) {
    p.T_w2c[tid] = sT[tid];
    if (p.viewmatrix) p.viewmatrix[tid] = sView[tid];
    if (p.projmatrix) p.projmatrix[tid] = sProj[tid];
    if (p.gather_out) p.gather_out[tid] = sT[tid];
  }
  if (tid < 8) p.params[tid] = sP[tid];
  if (p.gather_out && tid >= 16 && tid < 20) p.gather_out[tid] = tid == 16 ? sP[6] : (tid == 17 ? sP[7] : 0.f);
}

}  // namespace mgs

using namespace mgs;

static int check_map_args(const mgs_map_args* a) {
  if (!a) { set_error("map: args is null"); return MGS_EINVAL; }
  const mgs_gaussians& g = a->g;
  if (g.P < 0 || g.sh_coeffs < 1 || g.sh_degree < 0 || g.sh_degree > 3 || (g.sh_degree + 1) * (g.sh_degree + 1) > g.sh_coeffs) {
    set_error("map: bad P / sh_coeffs / sh_degree"); return MGS_EINVAL;
  }
  if (g.P > 0 && (!g.xyz || !g.f_dc || !g.opacity_raw || !g.scaling_raw || !g.rotation_raw || (g.sh_coeffs > 1 && !g.f_rest))) {
    set_error("map: null parameter pointer"); return MGS_EINVAL;
  }
  if (g.P > 0 && (!a->grad_xyz || !a->grad_f_dc || !a->grad_opacity || !a->grad_scaling || !a->grad_rotation ||
                  (g.sh_coeffs > 1 && !a->grad_f_rest))) { set_error("map: null gradient pointer"); return MGS_EINVAL; }
  if (!a->loss_total || !a->scratch) { set_error("map: loss_total / scratch is null"); return MGS_EINVAL; }
  return MGS_OK;
}

extern "C" {

size_t mgs_map_scratch_bytes(int32_t P, int32_t sh_coeffs, int32_t W_max, int32_t H_max) {
  if (P < 0 || sh_coeffs < 1 || W_max <= 0 || H_max <= 0) return 0;
  return carve_map_scratch(nullptr, P, sh_coeffs, W_max, H_max).total;
}

int mgs_map_begin(const mgs_map_args* a, void* stream) {
  int rc = check_map_args(a);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int P = a->g.P;
  // sizes of the image-dependent tail do not matter here: only the per-Gaussian head is touched
  const MapScratch m = carve_map_scratch(a->scratch, P, a->g.sh_coeffs, 16, 16);
  if (a->scratch_bytes < m.total) { set_error("map: scratch too small"); return MGS_ENOSPC; }
  if (P == 0) { MGS_CUDA_TRY(memset_async(a->loss_total, 0, sizeof(float), s)); return MGS_OK; }
  MGS_PROF("map_activate", s, map_activate_kernel<<<(P + 255) / 256, 256, 0, s>>>(a->g, m, a->grad_xyz, a->grad_f_dc, a->loss_total));
  MGS_LAUNCH_CHECK();
  const int M = a->g.sh_coeffs;
  if (M > 1) {   // SH gradients accumulate straight into the raw-parameter gradients (features have no activation)
    MGS_CUDA_TRY(memset_async(a->grad_f_dc, 0, (size_t)P * 3 * sizeof(float), s));
    MGS_CUDA_TRY(memset_async(a->grad_f_rest, 0, (size_t)P * (M - 1) * 3 * sizeof(float), s));
  }
  return MGS_OK;
}

// shared validation + derived state of the two halves of a view
struct MapViewCtx { ViewConst v; WsLayout L; MapScratch m; const float* shs; int P, M, tiles, use_depth, slot; size_t npix; };

static int map_view_ctx(const mgs_map_args* a, const mgs_map_view* mv, MapViewCtx* c) {
  int rc = check_map_args(a);
  if (rc) return rc;
  if (!mv) { set_error("map: view is null"); return MGS_EINVAL; }
  const mgs_view* view = &mv->view;
  if (view->image_height <= 0 || view->image_width <= 0 || !view->bg || !view->viewmatrix || !view->projmatrix ||
      !view->projmatrix_raw) { set_error("map: incomplete view"); return MGS_EINVAL; }
  if (!mv->gt_image || !mv->out_color || !mv->out_depth || !mv->out_opacity || !mv->radii || !mv->n_touched || !mv->workspace) {
    set_error("map: null view buffer"); return MGS_EINVAL;
  }
  if ((mv->exposure_a == nullptr) != (mv->exposure_b == nullptr)) { set_error("map: give both exposure parameters or neither"); return MGS_EINVAL; }
  if (mv->R_cap <= 0 || mv->R_cap > 0x7fffffffLL) { set_error("map: bad R_cap"); return MGS_EINVAL; }
  if (mv->slot < 0 || mv->slot >= MAP_SLOTS) { set_error("map: slot must be in [0, %d)", MAP_SLOTS); return MGS_EINVAL; }
  const mgs_gaussians& g = a->g;
  c->P = g.P; c->M = g.sh_coeffs; c->slot = mv->slot;
  const int W = view->image_width, H = view->image_height;
  if (c->P == 0) return MGS_OK;
  c->m = carve_map_scratch(a->scratch, c->P, c->M, W, H);
  if (a->scratch_bytes < c->m.total) { set_error("map: scratch too small for this view: %zu < %zu", a->scratch_bytes, c->m.total); return MGS_ENOSPC; }
  if ((rc = ws_layout(c->P, W, H, mv->R_cap, &c->L))) return rc;
  if (mv->workspace_bytes < c->L.total) { set_error("map: workspace too small: %zu < %zu", mv->workspace_bytes, c->L.total); return MGS_ENOSPC; }
  mgs_view vv = *view;
  vv.sh_degree = g.sh_degree; vv.sh_coeffs = c->M;
  if (g.sh_degree > 0 && !vv.campos) { set_error("map: campos required for sh_degree > 0"); return MGS_EINVAL; }
  c->v = make_view_const(&vv);
  c->npix = (size_t)W * H;
  c->tiles = c->v.gx * c->v.gy;
  c->use_depth = mv->gt_depth != nullptr;
  c->shs = g.f_dc;   // M > 1: split layout, f_rest passed beside it
  return MGS_OK;
}

int mgs_map_view_forward(const mgs_map_args* a, const mgs_map_view* mv, void* stream) {
  MapViewCtx c;
  int rc = map_view_ctx(a, mv, &c);
  if (rc) return rc;
  if (c.P == 0) return MGS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  PdlScope pdl_scope;    // preprocess .. loss finalize chained by programmatic dependent launch (common.cuh)
  const mgs_gaussians& g = a->g;
  const MapScratch& m = c.m;
  void* ws = mv->workspace;
  // n_touched is written by the forward blend only: cleared ahead of the chain so that no memset sits between its kernels
  MGS_CUDA_TRY(memset_async(mv->n_touched, 0, (size_t)c.P * sizeof(int32_t), s));
  // the workspace of a mapping engine is zero-filled once by its owner: the scan runs in the last CTA of the preprocess
  // kernel and the sort kernel leaves the tile counters zero for the next view (raster_forward.cu, FoldScan)
  // tile_segments: fixed 1024-slot segment per tile, scattered by the preprocess kernel (no scan, no emit, one sort kernel)
  const bool segments = mv->tile_segments != 0;
  if (segments && (int64_t)c.tiles * TILE_SEGMENT > mv->R_cap) { set_error("map: tile_segments needs R_cap >= tiles x %d", TILE_SEGMENT); return MGS_ENOSPC; }
  const FoldScan fs{mv->R_cap, /*heavy_first=*/true, /*in_kernel=*/!segments, segments ? TILE_SEGMENT : 0};
  if ((rc = launch_preprocess(c.v, c.P, g.xyz, m.act_opacity, c.shs, nullptr, m.act_scales, m.act_rot, nullptr, mv->radii, ws, c.L, s,
                              nullptr, 0, c.M > 1 ? g.f_rest : nullptr, &fs))) return rc;
  if (segments) { if ((rc = launch_sort_segments(c.v, ws, c.L, mv->R_cap, s, /*ids_only=*/false))) return rc; }
  else if ((rc = launch_binning(c.v, c.P, nullptr, mv->radii, ws, c.L, mv->R_cap, s, true, /*scan_done=*/true, /*clear_counts=*/true))) return rc;

  LossK lk;
  lk.mode = MGS_LOSS_MAP_L1; lk.npix = (int)c.npix; lk.use_depth = c.use_depth;
  lk.thr = mv->rgb_boundary_threshold; lk.alpha = mv->alpha; lk.delta = 0.f; lk.scale = 1.f; lk.eps = 1e-8f;
  lk.image = nullptr; lk.depth = nullptr; lk.opacity = nullptr; lk.gt = mv->gt_image; lk.gt_depth = mv->gt_depth;
  lk.mask = nullptr; lk.a = mv->exposure_a; lk.b = mv->exposure_b;
  FusedLoss fl;
  fl.k = &lk; fl.grad_scale = (c.use_depth ? mv->alpha : 1.f) / (3.f * (float)c.npix);
  fl.g_img = m.g_img[c.slot]; fl.g_depth = m.g_depth[c.slot]; fl.partials = m.partials[c.slot]; fl.fast = mv->fast_math != 0;
  if ((rc = launch_blend_forward(c.v, ws, c.L, mv->out_color, mv->out_depth, mv->out_opacity, mv->n_touched, s, &fl, /*heavy_first=*/true, /*clear_counts=*/segments))) return rc;
  cudaError_t e;
  MGS_PROF("map_loss_finalize", s, e = launch_chain(map_loss_finalize_kernel, 1, 256, 0, s, m.partials[c.slot], c.tiles, (int)c.npix, c.use_depth,
                                                    mv->alpha, mv->loss, a->loss_total, mv->dab));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

static int map_view_backward(const mgs_map_args* a, const mgs_map_view* mv, void* stream, bool pre_cleared) {
  MapViewCtx c;
  int rc = map_view_ctx(a, mv, &c);
  if (rc) return rc;
  if (c.P == 0) return MGS_OK;
  PdlScope pdl_scope;
  const mgs_gaussians& g = a->g;
  const MapScratch& m = c.m;
  BackwardArgs b;
  memset(&b, 0, sizeof(b));
  b.v = c.v; b.P = c.P;
  b.grad_mask = MGS_GRAD_MEANS3D | MGS_GRAD_COLOR | MGS_GRAD_OPACITY | MGS_GRAD_COV | (mv->tau ? MGS_GRAD_TAU : 0u);
  b.means3D = g.xyz; b.shs = c.shs; b.scales = m.act_scales; b.rotations = m.act_rot; b.radii = mv->radii;
  b.ws = mv->workspace; b.L = c.L; b.scratch = m.bwd[c.slot]; b.scratch_bytes = m.bwd_bytes;
  b.dL_dcolor_img = m.g_img[c.slot]; b.dL_ddepth_img = m.g_depth[c.slot];
  b.dL_dmeans3D = a->grad_xyz; b.dL_dcolors_or_sh = a->grad_f_dc; b.dL_dopacity = m.d_opacity;
  if (c.M > 1) { b.sh_rest = g.f_rest; b.dL_dsh_rest = a->grad_f_rest; }
  b.dL_dscales = m.d_scales; b.dL_drotations = m.d_rot; b.dL_dtau_sum = mv->tau;
  b.accumulate = mv->concurrent_backward ? 2 : 1;
  if (mv->update_stats) { b.stat_grad_accum = a->xyz_gradient_accum; b.stat_denom = a->denom; b.stat_max_radii = a->max_radii2D; }
  b.fast = mv->fast_math != 0;
  b.heavy_first = true;     // the order the view's scan left in its workspace
  b.pre_cleared = pre_cleared;
  return launch_backward_full(b, (cudaStream_t)stream);
}

int mgs_map_view_backward(const mgs_map_args* a, const mgs_map_view* mv, void* stream) { return map_view_backward(a, mv, stream, false); }

// Forward and backward of one view back to back on one stream (a rank that owns a single view of the window: the whole
// iteration is this dependent chain).  The backward's accumulators are cleared AHEAD of the forward, so the view is eight
// kernels with nothing between them, chained by programmatic dependent launch.
int mgs_map_view_fwd_bwd(const mgs_map_args* a, const mgs_map_view* mv, void* stream) {
  MapViewCtx c;
  int rc = map_view_ctx(a, mv, &c);
  if (rc) return rc;
  if (c.P == 0) return MGS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  PdlScope pdl_scope;
  MGS_CUDA_TRY(memset_async(c.m.bwd[c.slot], 0, (size_t)c.P * backward_g2d_bytes(), s));
  if (mv->tau) MGS_CUDA_TRY(memset_async(mv->tau, 0, 6 * sizeof(float), s));
  rc = mgs_map_view_forward(a, mv, stream);
  return rc ? rc : map_view_backward(a, mv, stream, true);
}

int mgs_window_pose_step(const mgs_pose_slot* slots, int32_t n, const float* proj_raw, float beta1, float beta2, float eps,
                         void* stream) {
  if (n == 0) return MGS_OK;
  if (!slots || n < 0 || n > POSE_SLOTS_MAX || !proj_raw) { set_error("window_pose_step: 1..%d slots and a projection matrix", POSE_SLOTS_MAX); return MGS_EINVAL; }
  PoseSlots k;
  memset(&k, 0, sizeof(k));
  for (int i = 0; i < n; i++) {
    const mgs_pose_slot& p = slots[i];
    if (!p.T_w2c || !p.params || (!p.frozen && (!p.exp_avg || !p.exp_avg_sq || !p.step || !p.tau || !p.dab))) {
      set_error("window_pose_step: slot %d has null pointers", i); return MGS_EINVAL;
    }
    k.s[i] = p;
  }
  k.proj_raw = proj_raw; k.beta1 = beta1; k.beta2 = beta2; k.eps = eps;
  cudaStream_t s = (cudaStream_t)stream;
  MGS_PROF("window_pose_step", s, window_pose_step_kernel<<<n, 32, 0, s>>>(k));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_map_end(const mgs_map_args* a, int32_t add_regulariser, float regulariser_weight, void* stream) {
  int rc = check_map_args(a);
  if (rc) return rc;
  const int P = a->g.P;
  if (P == 0) return MGS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const MapScratch m = carve_map_scratch(a->scratch, P, a->g.sh_coeffs, 16, 16);
  MGS_PROF("map_end", s, map_end_kernel<<<(P + 255) / 256, 256, 0, s>>>(a->g, m, a->grad_f_dc, a->grad_f_rest, a->grad_opacity,
                                                                       a->grad_scaling, a->grad_rotation,
                                                                       add_regulariser ? regulariser_weight : 0.f, a->loss_total));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
