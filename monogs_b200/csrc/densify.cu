// Densify / clone / split / prune of the Gaussian parameter arena with its Adam state, in two launches
// (SURVEY.md §8 row f3).  Replaces GaussianModel.densify_and_prune / densify_and_clone / densify_and_split /
// prune_points / cat_tensors_to_optimizer / _prune_optimizer (gaussian_splatting/scene/gaussian_model.py:470-691),
// which run dozens of boolean-index / cat kernels and rebuild every nn.Parameter and optimizer state three times
// (clone, split, prune).
//
// Everything the reference decides is a per-SOURCE-Gaussian function, so one pass plans and one pass moves:
//   plan  : for Gaussian g  ->  keep (0/1), clone (0/1), children (0 or N), split-selected (0/1)
//           keep     = not split-selected and not pruned
//           clone    = grad >= thr and max(scale) <= percent_dense*extent, and not pruned (a clone carries g's values)
//           children = grad >= thr and max(scale) >  percent_dense*extent, unless the child itself is pruned
//                      (opacity < min or max(scale)/(0.8 N) > 0.1 extent)
//           pruned   = opacity < min_opacity or max(scale) > 0.1 extent  [or the caller's mask]
//           (the reference resets max_radii2D to zero in densification_postfix BEFORE its final prune, so its
//            max_screen_size test can never fire inside densify_and_prune: reproduced by not testing it)
//           + an exclusive scan of the four counters.
//   apply : the final arena is  [kept originals | clones | children rep 0 | ... | children rep N-1], each segment in
//           ascending source order — exactly the order the reference's cat + boolean-mask sequence produces.
//           Adam moments are copied for kept originals and zero for new points (cat_tensors_to_optimizer).
// Split samples: new_xyz = R(q) (noise * scale) + xyz with noise ~ N(0,1) supplied by the caller in the reference's
// own layout [N * n_split_selected, 3] (row r * n_sel + k), so the arithmetic is testable against a torch restatement;
// the draws themselves can not be pinned to torch's CUDA generator.
#include "launchers.h"
#include <cub/device/device_scan.cuh>

namespace mgs {

struct Cnt4 { uint32_t keep, clone, child, sel; };
struct Cnt4Add {
  __host__ __device__ __forceinline__ Cnt4 operator()(const Cnt4& a, const Cnt4& b) const {
    return Cnt4{a.keep + b.keep, a.clone + b.clone, a.child + b.child, a.sel + b.sel};
  }
};

__global__ void __launch_bounds__(256)
densify_plan_kernel(const int P, const int isotropic, const float* __restrict__ opacity_raw,
                    const float* __restrict__ scaling_raw, const float* __restrict__ grad_accum,
                    const float* __restrict__ denom, const uint8_t* __restrict__ prune_mask, const mgs_densify_cfg cfg,
                    Cnt4* __restrict__ cnt) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  Cnt4 c{0u, 0u, 0u, 0u};
  if (cfg.mode == 1) {   // prune_points(mask) only
    c.keep = prune_mask[idx] ? 0u : 1u;
  } else {
    float grad = 0.f;
    if (grad_accum != nullptr) { grad = grad_accum[idx] / denom[idx]; if (grad != grad) grad = 0.f; }   // grads[grads.isnan()] = 0
    float smax;
    if (isotropic) smax = expf(scaling_raw[idx]);
    else smax = fmaxf(fmaxf(expf(scaling_raw[3 * idx]), expf(scaling_raw[3 * idx + 1])), expf(scaling_raw[3 * idx + 2]));
    const float op = 1.f / (1.f + expf(-opacity_raw[idx]));
    const bool hot = grad >= cfg.grad_threshold;
    const bool small = smax <= cfg.percent_dense * cfg.extent;
    const bool low = op < cfg.min_opacity;
    bool pruned = low || (cfg.max_screen_size > 0.f && smax > 0.1f * cfg.extent);
    if (prune_mask != nullptr && prune_mask[idx]) pruned = true;
    const bool sel = hot && !small;
    c.sel = sel ? 1u : 0u;
    c.keep = (!sel && !pruned) ? 1u : 0u;
    c.clone = (hot && small && !pruned) ? 1u : 0u;
    if (sel) {
      const float child_smax = smax / (0.8f * (float)cfg.n_split);
      bool child_pruned = low || (cfg.max_screen_size > 0.f && child_smax > 0.1f * cfg.extent);
      if (prune_mask != nullptr && prune_mask[idx]) child_pruned = true;
      c.child = child_pruned ? 0u : 1u;
    }
  }
  cnt[idx] = c;
}

__global__ void densify_totals_kernel(const Cnt4* __restrict__ cnt, const Cnt4* __restrict__ excl, const int P,
                                      int32_t* __restrict__ totals /* keep, clone, children(per rep), split-selected */) {
  const Cnt4 a = excl[P - 1], b = cnt[P - 1];
  totals[0] = (int32_t)(a.keep + b.keep); totals[1] = (int32_t)(a.clone + b.clone);
  totals[2] = (int32_t)(a.child + b.child); totals[3] = (int32_t)(a.sel + b.sel);
}

struct ArenaPtrs {
  const float* src[6]; const float* src_m[6]; const float* src_v[6];
  float* dst[6]; float* dst_m[6]; float* dst_v[6];
  int width[6];
  const int32_t *src_kf, *src_nobs; int32_t *dst_kf, *dst_nobs;
};

__device__ __forceinline__ void copy_row(const ArenaPtrs& A, const int a, const size_t from, const size_t to, const bool state) {
  const int w = A.width[a];
  if (w == 0 || A.src[a] == nullptr) return;
  for (int k = 0; k < w; k++) A.dst[a][to * w + k] = A.src[a][from * w + k];
  if (A.dst_m[a] != nullptr)
    for (int k = 0; k < w; k++) A.dst_m[a][to * w + k] = (state && A.src_m[a]) ? A.src_m[a][from * w + k] : 0.f;
  if (A.dst_v[a] != nullptr)
    for (int k = 0; k < w; k++) A.dst_v[a][to * w + k] = (state && A.src_v[a]) ? A.src_v[a][from * w + k] : 0.f;
}

__global__ void __launch_bounds__(256)
densify_apply_kernel(const int P, const int isotropic, const int n_split, const Cnt4* __restrict__ cnt,
                     const Cnt4* __restrict__ excl, const int32_t* __restrict__ totals, const ArenaPtrs A,
                     const float* __restrict__ noise) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const Cnt4 c = cnt[idx], e = excl[idx];
  const size_t n_keep = (size_t)totals[0], n_clone = (size_t)totals[1], n_child = (size_t)totals[2], n_sel = (size_t)totals[3];
  auto bookkeeping = [&](size_t to) {
    if (A.dst_kf != nullptr) A.dst_kf[to] = A.src_kf[idx];
    if (A.dst_nobs != nullptr) A.dst_nobs[to] = A.src_nobs[idx];
  };
  if (c.keep) {
    const size_t to = e.keep;
    for (int a = 0; a < 6; a++) copy_row(A, a, idx, to, true);
    bookkeeping(to);
  }
  if (c.clone) {
    const size_t to = n_keep + e.clone;
    for (int a = 0; a < 6; a++) copy_row(A, a, idx, to, false);
    bookkeeping(to);
  }
  if (c.child) {
    // gaussian_model.py:615-631: samples = N(0, scale); new_xyz = R(q) samples + xyz; new_scaling = log(scale / (0.8 N))
    const float4 q = reinterpret_cast<const float4*>(A.src[5])[idx];
    const float nrm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    const float r = q.x / nrm, x = q.y / nrm, y = q.z / nrm, z = q.w / nrm;
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float s[3];
    if (isotropic) s[0] = s[1] = s[2] = expf(A.src[4][idx]);
    else { s[0] = expf(A.src[4][3 * idx]); s[1] = expf(A.src[4][3 * idx + 1]); s[2] = expf(A.src[4][3 * idx + 2]); }
    const float3 p = {A.src[0][3 * idx], A.src[0][3 * idx + 1], A.src[0][3 * idx + 2]};
    for (int rep = 0; rep < n_split; rep++) {
      const size_t to = n_keep + n_clone + (size_t)rep * n_child + e.child;
      for (int a = 1; a < 6; a++) copy_row(A, a, idx, to, false);     // f_dc, f_rest, opacity, (scaling), rotation
      const float* nz = noise + ((size_t)rep * n_sel + e.sel) * 3;
      const float v[3] = {nz[0] * s[0], nz[1] * s[1], nz[2] * s[2]};
      A.dst[0][3 * to + 0] = R[0][0] * v[0] + R[0][1] * v[1] + R[0][2] * v[2] + p.x;
      A.dst[0][3 * to + 1] = R[1][0] * v[0] + R[1][1] * v[1] + R[1][2] * v[2] + p.y;
      A.dst[0][3 * to + 2] = R[2][0] * v[0] + R[2][1] * v[1] + R[2][2] * v[2] + p.z;
      if (A.dst_m[0]) { A.dst_m[0][3 * to] = 0.f; A.dst_m[0][3 * to + 1] = 0.f; A.dst_m[0][3 * to + 2] = 0.f; }
      if (A.dst_v[0]) { A.dst_v[0][3 * to] = 0.f; A.dst_v[0][3 * to + 1] = 0.f; A.dst_v[0][3 * to + 2] = 0.f; }
      const float div = 0.8f * (float)n_split;
      if (isotropic) A.dst[4][to] = logf(s[0] / div);
      else { A.dst[4][3 * to] = logf(s[0] / div); A.dst[4][3 * to + 1] = logf(s[1] / div); A.dst[4][3 * to + 2] = logf(s[2] / div); }
      bookkeeping(to);
    }
  }
}

static size_t densify_scan_bytes(int P) {
  size_t b = 0;
  cub::DeviceScan::ExclusiveScan(nullptr, b, (const Cnt4*)nullptr, (Cnt4*)nullptr, Cnt4Add(), Cnt4{0u, 0u, 0u, 0u}, P > 0 ? P : 1);
  return b;
}

}  // namespace mgs

using namespace mgs;

extern "C" {

size_t mgs_densify_scratch_bytes(int32_t P) {
  const size_t Pp = P > 0 ? P : 1;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  return up(Pp * sizeof(Cnt4)) * 2 + up(densify_scan_bytes(P)) + 256;
}

int mgs_densify_plan(const mgs_gaussians* g, const float* xyz_gradient_accum, const float* denom, const uint8_t* prune_mask,
                     const mgs_densify_cfg* cfg, void* scratch, size_t scratch_bytes, int32_t* totals_out, void* stream) {
  if (!g || !cfg || !scratch || !totals_out) { set_error("densify_plan: null argument"); return MGS_EINVAL; }
  if (g->P <= 0) { set_error("densify_plan: P must be positive"); return MGS_EINVAL; }
  if (cfg->mode == 1 && !prune_mask) { set_error("densify_plan: mode 1 needs prune_mask"); return MGS_EINVAL; }
  if (cfg->mode == 0 && (!g->opacity_raw || !g->scaling_raw || (xyz_gradient_accum && !denom))) { set_error("densify_plan: null input"); return MGS_EINVAL; }
  if (cfg->mode == 0 && cfg->n_split < 1) { set_error("densify_plan: n_split must be >= 1"); return MGS_EINVAL; }
  if (scratch_bytes < mgs_densify_scratch_bytes(g->P)) { set_error("densify_plan: scratch too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  const int P = g->P;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  char* sp = reinterpret_cast<char*>(scratch);
  Cnt4* cnt = reinterpret_cast<Cnt4*>(sp); sp += up((size_t)P * sizeof(Cnt4));
  Cnt4* excl = reinterpret_cast<Cnt4*>(sp); sp += up((size_t)P * sizeof(Cnt4));
  int32_t* totals = reinterpret_cast<int32_t*>(sp); sp += 256;
  size_t tmp = densify_scan_bytes(P);
  densify_plan_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g->isotropic, g->opacity_raw, g->scaling_raw, xyz_gradient_accum, denom,
                                                      prune_mask, *cfg, cnt);
  MGS_LAUNCH_CHECK();
  MGS_CUDA_TRY(cub::DeviceScan::ExclusiveScan(sp, tmp, (const Cnt4*)cnt, excl, Cnt4Add(), Cnt4{0u, 0u, 0u, 0u}, P, s));
  densify_totals_kernel<<<1, 1, 0, s>>>(cnt, excl, P, totals);
  MGS_LAUNCH_CHECK();
  MGS_CUDA_TRY(cudaMemcpyAsync(totals_out, totals, 4 * sizeof(int32_t), cudaMemcpyDefault, s));
  return MGS_OK;
}

int mgs_densify_apply(const mgs_gaussians* g, const mgs_gaussian_state* src, const mgs_gaussian_state* dst,
                      const mgs_densify_cfg* cfg, const float* noise, const void* scratch, size_t scratch_bytes, void* stream) {
  if (!g || !src || !dst || !cfg || !scratch) { set_error("densify_apply: null argument"); return MGS_EINVAL; }
  if (g->P <= 0) { set_error("densify_apply: P must be positive"); return MGS_EINVAL; }
  if (scratch_bytes < mgs_densify_scratch_bytes(g->P)) { set_error("densify_apply: scratch too small"); return MGS_ENOSPC; }
  const int P = g->P, M = g->sh_coeffs;
  ArenaPtrs A;
  const int width[6] = {3, 3, 3 * (M - 1), 1, g->isotropic ? 1 : 3, 4};
  for (int a = 0; a < 6; a++) {
    A.width[a] = width[a];
    A.src[a] = src->param[a]; A.src_m[a] = src->exp_avg[a]; A.src_v[a] = src->exp_avg_sq[a];
    A.dst[a] = dst->param[a]; A.dst_m[a] = dst->exp_avg[a]; A.dst_v[a] = dst->exp_avg_sq[a];
    if (width[a] > 0 && (!A.src[a] || !A.dst[a])) { set_error("densify_apply: parameter array %d is null", a); return MGS_EINVAL; }
  }
  A.src_kf = src->unique_kfIDs; A.src_nobs = src->n_obs;
  A.dst_kf = src->unique_kfIDs ? dst->unique_kfIDs : nullptr; A.dst_nobs = src->n_obs ? dst->n_obs : nullptr;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  const char* sp = reinterpret_cast<const char*>(scratch);
  const Cnt4* cnt = reinterpret_cast<const Cnt4*>(sp); sp += up((size_t)P * sizeof(Cnt4));
  const Cnt4* excl = reinterpret_cast<const Cnt4*>(sp); sp += up((size_t)P * sizeof(Cnt4));
  const int32_t* totals = reinterpret_cast<const int32_t*>(sp);
  cudaStream_t s = (cudaStream_t)stream;
  densify_apply_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g->isotropic, cfg->mode == 0 ? cfg->n_split : 0, cnt, excl, totals, A, noise);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
